"""GPU: the allocation-free scatter entry point (tcnn_b200_encode_backward_noalloc, include/tcnn_b200.h) against the
existing tcnn_b200_encode_backward, through the C ABI, on two calls in a row.

The new entry point keeps its scratch (the private gradient copies of the coarse levels) in a file-scope device array that
every call shares and that keeps the previous call's contents; the second call, on a different gradient, shows that the
first one's leftovers are cleared.  Tolerance: the two entry points issue the same fp16 reductions, whose order is not
deterministic - 1e-2 relative L2, the bar test_gpu_parity.py uses for grid gradients ("fp16 red accumulation order").

Named test_zz_* so that pytest collects it last."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from tests import common  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _scratch_limit():
    header = open(os.path.join(ROOT, "include", "tcnn_b200.h")).read()
    m = re.search(r"#define\s+TCNN_B200_NOALLOC_SCRATCH_BYTES\s+\((\d+)u\s*<<\s*(\d+)\)", header)
    assert m, "TCNN_B200_NOALLOC_SCRATCH_BYTES not found in include/tcnn_b200.h"
    return int(m.group(1)) << int(m.group(2))


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def test_noalloc_scatter_matches_workspace_scatter_twice():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device: the gpu-marked tests need a B200 (there is no CPU fallback)")
    import tinycudann as tcnn
    from tinycudann._lib import check, ptr

    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    native = tcnn.NativeModule(2, 3, common.CONFIG_1)
    lib, info = native.lib, native.info
    B = 1 << 15  # >= 2^14: the coarse levels scatter into the replicas, i.e. the scratch is really used
    E, n_grid = info.encoded_width, int(info.n_params_grid)
    ws = native.workspace(B, dev)
    gen = torch.Generator(device=dev).manual_seed(20)
    for call in range(2):
        x = torch.rand(B, 2, device=dev, generator=gen)
        dy = (torch.randn(B, E, device=dev, generator=gen) * (0.05 * (call + 1))).half()
        want = torch.full((n_grid,), 7.0, dtype=torch.half, device=dev)  # zero_first = 1 must clear the buffer
        got = torch.full((n_grid,), -3.0, dtype=torch.half, device=dev)
        check(lib.tcnn_b200_encode_backward(native.handle, _stream(), B, ptr(x), ptr(dy), ptr(want), 1, ptr(ws)))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        check(lib.tcnn_b200_encode_backward_noalloc(native.handle, _stream(), B, ptr(x), ptr(dy), ptr(got), 1))
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(dev)[0] >= free_before, "the allocation-free call took device memory"
        w, g = want.float().cpu().numpy(), got.float().cpu().numpy()
        assert np.isfinite(g).all()
        assert np.abs(w).max() > 0
        for l in range(info.n_levels):  # per level: the replicated coarse levels must not hide behind the large fine ones
            a, b = 2 * info.level_offset[l], 2 * info.level_offset[l + 1]
            assert common.rel_l2(g[a:b], w[a:b]) < 1e-2, f"call {call}, level {l}"
        assert common.rel_l2(g, w) < 1e-2, f"call {call}"


def test_noalloc_rejects_grids_beyond_the_scratch_limit():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device: the gpu-marked tests need a B200 (there is no CPU fallback)")
    import tinycudann as tcnn
    from tinycudann._lib import ptr

    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    # 32 levels of <= 8192 entries, 8 features: 32 replicas of every level = far more than the built-in array
    cfg = json.loads(json.dumps(common.CONFIG_1))
    cfg["encoding"] = {"otype": "HashGrid", "n_levels": 32, "n_features_per_level": 4, "log2_hashmap_size": 12, "base_resolution": 16, "per_level_scale": 1.2}
    native = tcnn.NativeModule(2, 3, cfg)
    info = native.info
    B = 1 << 14
    sizes = [info.level_offset[l + 1] - info.level_offset[l] for l in range(info.n_levels)]
    need = sum(s for s in sizes if s <= 8192) * info.n_features_per_level * 2 * 32  # 32 fp16 copies of every coarse level
    assert need > _scratch_limit()
    x = torch.rand(B, 2, device=dev)
    dy = torch.zeros(B, info.encoded_width, dtype=torch.half, device=dev)
    grads = torch.zeros(int(info.n_params_grid), dtype=torch.half, device=dev)
    rc = native.lib.tcnn_b200_encode_backward_noalloc(native.handle, _stream(), B, ptr(x), ptr(dy), ptr(grads), 1)
    assert rc != 0
    msg = native.lib.tcnn_b200_last_error().decode()
    assert "scratch" in msg and str(_scratch_limit()) in msg, msg
    torch.cuda.synchronize()
