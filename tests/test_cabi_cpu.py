"""CPU: the C-ABI library loads, exports every symbol include/tcnn_b200.h declares, resolves configs exactly like the
reference's factories, reports errors through tcnn_b200_last_error, and refuses to compute without a GPU."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

from tests import common

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge
    ge.build_library()
    from tinycudann import _lib
    return _lib.lib()


def _plan(lib, cfg, n_in, n_out):
    from tinycudann import _lib
    h = C.c_void_p()
    rc = lib.tcnn_b200_plan_create(json.dumps(cfg).encode(), n_in, n_out, C.byref(h))
    return rc, h


def test_exports_every_declared_symbol(lib):
    from tinycudann import _lib
    header = open(os.path.join(ROOT, "include", "tcnn_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(tcnn_b200_[a-z0-9_]+)\s*\(", header)))
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in tcnn_b200.h but not exported"
    assert set(declared) == set(_lib.exported_symbols()), "ctypes binding out of sync with the header"
    assert lib.tcnn_b200_arch() == b"sm_100a"
    assert lib.tcnn_b200_batch_size_granularity() == 256


@pytest.mark.parametrize("cfg,n_in,n_out,n_grid", [
    (common.CONFIG_1, 2, 3, 11184640),
    (common.CONFIG_SHIPPED, 2, 3, 708368),
    (common.CONFIG_NERF, 3, 4, 12196240),
])
def test_plan_sizes_match_reference_tables(lib, cfg, n_in, n_out, n_grid):
    from tinycudann import _lib
    rc, h = _plan(lib, cfg, n_in, n_out)
    assert rc == 0, lib.tcnn_b200_last_error()
    info = _lib.Info()
    assert lib.tcnn_b200_plan_info(h, C.byref(info)) == 0
    assert info.n_params_grid == n_grid
    assert info.n_params_mlp == 64 * 32 + 64 * 64 + 16 * 64
    assert info.encoded_width == 32 and info.padded_output_width == 16
    g, m = common.oracle_configs(cfg, n_in, n_out)
    assert np.array_equal(np.array(info.level_offset[:17], dtype=np.uint32), g.offsets)
    assert lib.tcnn_b200_workspace_bytes(h, 1 << 18) > 2 * (1 << 18) * 32 * 2
    lib.tcnn_b200_plan_destroy(h)


def test_defaults_and_case_insensitive_otype(lib):
    from tinycudann import _lib
    cfg = {"encoding": {"otype": "hashgrid"}, "network": {"otype": "FULLYFUSEDMLP", "n_neurons": 64, "n_hidden_layers": 2}}
    rc, h = _plan(lib, cfg, 3, 1)
    assert rc == 0, lib.tcnn_b200_last_error()
    info = _lib.Info()
    lib.tcnn_b200_plan_info(h, C.byref(info))
    assert (info.n_levels, info.n_features_per_level, info.log2_hashmap_size, info.base_resolution) == (16, 2, 19, 16)
    assert info.per_level_scale == 2.0 and info.loss_type == 1  # default loss is RelativeL2 (src/loss.cu:86)
    assert abs(info.learning_rate - 1e-3) < 1e-9 and abs(info.beta2 - 0.999) < 1e-7 and abs(info.epsilon - 1e-8) < 1e-12
    assert lib.tcnn_b200_plan_update_optimizer(h, json.dumps({"learning_rate": 0.01, "beta2": 0.99}).encode()) == 0
    lib.tcnn_b200_plan_info(h, C.byref(info))
    assert abs(info.learning_rate - 0.01) < 1e-8 and abs(info.beta2 - 0.99) < 1e-7
    lib.tcnn_b200_plan_destroy(h)


@pytest.mark.parametrize("patch,msg", [
    ({"encoding": {"otype": "Frequency"}}, "outside the hot path"),
    ({"network": {"otype": "CutlassMLP"}}, "outside the hot path"),
    ({"network": {"otype": "FullyFusedMLP", "n_neurons": 48}}, "n_neurons"),
    ({"encoding": {"otype": "HashGrid", "n_levels": 4, "n_features": 8}}, "simultaneously"),
    ({"loss": {"otype": "CrossEntropy"}}, "outside the hot path"),
    ({"optimizer": {"otype": "Shampoo"}}, "outside the hot path"),
    ({"optimizer": {"otype": "Adam", "learnign_rate": 1e-2}}, "unknown optimizer key"),          # a misspelt key must not train with the default
    ({"encoding": {"otype": "TiledGrid", "n_levels": 4, "base_resolution": 5}}, "multiple of 8"),  # 25-entry tiles: unaligned level tables
])
def test_error_reporting(lib, patch, msg):
    cfg = json.loads(json.dumps(common.CONFIG_1))
    cfg.update(patch)
    rc, _ = _plan(lib, cfg, 2, 3)
    assert rc != 0
    assert msg in lib.tcnn_b200_last_error().decode()


def test_optimizer_freeze_flags_are_accepted(lib):
    """optimize_matrix_params / optimize_non_matrix_params (optimizers/adam.h:57-58, 77-85) are keys of the reference's Adam."""
    cfg = json.loads(json.dumps(common.CONFIG_1))
    cfg["optimizer"] = {"otype": "Adam", "optimize_matrix_params": False, "optimize_non_matrix_params": True, "adabound": False}
    rc, h = _plan(lib, cfg, 2, 3)
    assert rc == 0, lib.tcnn_b200_last_error()
    assert lib.tcnn_b200_plan_update_optimizer(h, json.dumps({"optimize_non_matrix_params": False}).encode()) == 0
    assert lib.tcnn_b200_plan_update_optimizer(h, json.dumps({"momentum": 0.9}).encode()) != 0
    lib.tcnn_b200_plan_destroy(h)


def test_malformed_json(lib):
    h = C.c_void_p()
    assert lib.tcnn_b200_plan_create(b'{"encoding": {', 2, 3, C.byref(h)) != 0
    assert b"json" in lib.tcnn_b200_last_error()


def test_no_cpu_fallback(lib):
    """Without a CUDA device every compute entry point must fail loudly (the product has no CPU path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; covered by the gpu tests")
    rc, h = _plan(lib, common.CONFIG_1, 2, 3)
    assert rc == 0
    buf = (C.c_float * 1024)()
    rc = lib.tcnn_b200_encode_forward(h, None, 256, buf, buf, buf)
    assert rc != 0 and b"no CUDA device" in lib.tcnn_b200_last_error()
    rc = lib.tcnn_b200_adam_step(h, None, 128.0, 1, buf, buf, buf, buf, buf, buf, 1)
    assert rc != 0
    lib.tcnn_b200_plan_destroy(h)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "tiny-cuda-nn_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(d, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle|#include\s+[\"<][^\n]*oracle", src, re.M), f


def test_cpp_facade_compiles_and_builds_a_plan_without_gpu():
    """include/tiny-cuda-nn/*.h (create_from_config / Trainer / Network / GPUMatrix / cpp::Module over the C ABI) compile
    with plain g++; without a device the sample host can still build the plan and report the reference's sizes."""
    import subprocess
    import __graft_entry__ as ge
    exe = ge.build_cpp_facade_sample()
    r = subprocess.run([exe, "--no-gpu"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "n_params=11191808 padded_output_width=16" in r.stdout


def test_snapshot_format_roundtrip_through_msgpack():
    """checkpoint.pack/unpack = the reference's Trainer::serialize object (trainer.h:275-315, adam.h:278-299); through msgpack
    (what hosts of the reference write with json::to_msgpack) and through nlohmann's text form of binary values."""
    import msgpack
    import numpy as np
    from tinycudann import checkpoint
    rng = np.random.default_rng(0)
    p = rng.standard_normal(1000).astype(np.float16)
    opt = {"current_step": 17, "base_learning_rate": 1e-2, "first_moments": rng.standard_normal(1000).astype(np.float32),
           "second_moments": rng.random(1000).astype(np.float32), "param_steps": rng.integers(0, 17, 1000).astype(np.uint32)}
    snap = checkpoint.pack(p, opt)
    assert snap["params_type"] == "__half" and snap["n_params"] == 1000 and len(snap["params_binary"]) == 2000
    back = msgpack.unpackb(msgpack.packb(snap, use_bin_type=True), raw=False)
    p2, o2 = checkpoint.unpack(back)
    assert p2.dtype == np.float16 and np.array_equal(p2, p)
    assert o2["current_step"] == 17 and abs(o2["base_learning_rate"] - 1e-2) < 1e-12
    for k in ("first_moments", "second_moments", "param_steps"):
        assert np.array_equal(o2[k], opt[k])
    # float snapshot, JSON-text binary form, missing param_steps (older snapshots: adam.h:290-295)
    f32 = {"n_params": 4, "params_type": "float", "params_binary": {"bytes": list(np.arange(4, dtype=np.float32).tobytes()), "subtype": None},
           "optimizer": {"current_step": 1, "base_learning_rate": 0.5, "first_moments_binary": np.zeros(4, np.float32).tobytes(),
                         "second_moments_binary": np.ones(4, np.float32).tobytes()}}
    p3, o3 = checkpoint.unpack(f32)
    assert p3.dtype == np.float32 and np.array_equal(p3, np.arange(4, dtype=np.float32)) and not o3["param_steps"].any()
    import pytest
    with pytest.raises(RuntimeError, match="float of __half"):
        checkpoint.unpack({"params_type": "double", "params_binary": b""})


def test_snapshot_written_by_the_reference_is_read():
    """tests/golden/snapshot_small_2d.msgpack = json::to_msgpack(Trainer::serialize(true)) of the REFERENCE build after 20 training
    steps on common.CONFIG_SNAPSHOT (`ref_driver serialize`, tests/dev/n2_pin.py).  checkpoint.unpack reads it; its sizes are
    the plan's; the C++ json type of include/tiny-cuda-nn/json.h re-emits the file byte for byte."""
    import json as js
    import subprocess
    import msgpack
    import numpy as np
    import __graft_entry__ as ge
    from tinycudann import checkpoint
    from tinycudann import _lib
    golden = os.path.join(ROOT, "tests", "golden")
    meta = js.load(open(os.path.join(golden, "snapshot_small_2d.json")))
    raw = open(os.path.join(golden, "snapshot_small_2d.msgpack"), "rb").read()
    params, opt = checkpoint.unpack(msgpack.unpackb(raw, raw=False))
    lib = _lib.lib()
    rc, h = _plan(lib, meta["config"], meta["n_in"], meta["n_out"])
    assert rc == 0, lib.tcnn_b200_last_error()
    info = _lib.Info()
    assert lib.tcnn_b200_plan_info(h, C.byref(info)) == 0
    P = int(info.n_params_grid + info.n_params_mlp)
    lib.tcnn_b200_plan_destroy(h)
    assert params.dtype == np.float16 and params.size == P
    assert opt["current_step"] == meta["n_steps"] and abs(opt["base_learning_rate"] - 1e-2) < 1e-9
    assert opt["first_moments"].size == P and opt["second_moments"].size == P and opt["param_steps"].size == P
    assert opt["param_steps"].max() == meta["n_steps"] and (opt["second_moments"] >= 0).all()
    # our writer -> the same object again
    again = msgpack.unpackb(msgpack.packb(checkpoint.pack(params, opt), use_bin_type=True), raw=False)
    p2, o2 = checkpoint.unpack(again)
    assert np.array_equal(p2, params) and all(np.array_equal(o2[k], opt[k]) for k in ("first_moments", "second_moments", "param_steps"))
    r = subprocess.run([ge.build_cpp_facade_sample(), "--msgpack", os.path.join(golden, "snapshot_small_2d.msgpack")], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and f"n_params={P} params_type=__half" in r.stdout and "rewritten_identically=1" in r.stdout, r.stdout + r.stderr


def test_network_widths(lib):
    """FullyFusedMLP widths of the reference (fully_fused_mlp.cu:862-870): 16, 32, 64, 128 - parameter counts W*E + (H-1)*W*W + 16*W."""
    from tinycudann import _lib
    for W, H in ((16, 2), (32, 3), (64, 2), (128, 3)):
        cfg = dict(common.CONFIG_1, network=dict(common.CONFIG_1["network"], n_neurons=W, n_hidden_layers=H))
        rc, h = _plan(lib, cfg, 2, 3)
        assert rc == 0, lib.tcnn_b200_last_error()
        info = _lib.Info()
        assert lib.tcnn_b200_plan_info(h, C.byref(info)) == 0
        assert info.n_params_mlp == W * 32 + (H - 1) * W * W + 16 * W
        lib.tcnn_b200_plan_destroy(h)
    rc, _ = _plan(lib, dict(common.CONFIG_1, network=dict(common.CONFIG_1["network"], n_neurons=48)), 2, 3)
    assert rc != 0 and b"16, 32, 64, and 128" in lib.tcnn_b200_last_error()
