"""CPU: pins the numpy oracle to the reference.  The fixtures in tests/golden/ are dumps of the reference's own CUDA
build (oracle/_ref/ref_driver over the unmodified sources) on a B200, made by tests/golden/make_golden.py."""
import json
import os

import numpy as np
import pytest

from oracle import tcnn_oracle as O
from tests import common

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
NAMES = ["cfg1_2d_T19", "shipped_2d_T15", "nerf_3d_T19", "f4_3d_T14", "tiled_smooth_2d", "dense_3d_f4", "nearest_3d_T12", "maxlevel_f4_3d_T14", "stochastic_2d_T12",
         "f1_2d_T12", "hash_prime_3d_T12", "hash_reversed_2d_T12", "hash_rng_3d_T12", "maxlevelgpu_f4_3d_T14", "accumulate_2d_T12",
         "narrow32_2d_T12", "narrow16_3d_T12"]


@pytest.fixture(scope="module", params=NAMES)
def case(request):
    z = np.load(os.path.join(GOLDEN, request.param + ".npz"))
    cfg = json.loads(str(z["config"]))
    n_in, n_out, B = int(z["n_in"]), int(z["n_out"]), int(z["B"])
    g, m = common.oracle_configs(cfg, n_in, n_out, scales=z["level_scale"])
    g.max_level = common.MAX_LEVEL_CASES.get(request.param, g.max_level)  # GridEncoding::set_max_level of that dump
    if request.param in common.MAX_LEVEL_GPU_CASES:                       # GridEncoding::set_max_level_gpu of that dump
        g.max_level_gpu = common.max_level_per_sample(B)
    x, t = common.synth_inputs(B, n_in, n_out, seed=int(z["seed"]))
    p = O.init_params(m, g, 1337, True)
    return dict(z=z, cfg=cfg, g=g, m=m, x=x, t=t, p=p, n_out=n_out, B=B, name=request.param)


def test_level_table(case):
    z, g = case["z"], case["g"]
    assert np.array_equal(g.resolutions, z["level_res"])
    assert np.array_equal(np.diff(g.offsets.astype(np.int64)), z["level_size"])
    assert g.n_params + case["m"].n_params == int(z["n_params"])


def test_param_init_bit_exact(case):
    z, p, m = case["z"], case["p"], case["m"]
    assert np.array_equal(p[: m.n_params], z["params_init_mlp"])
    assert np.array_equal(p[::997], z["params_init_sample"])
    assert p.astype(np.float64).sum() == float(z["params_init_sum"])
    assert (p.astype(np.float64) ** 2).sum() == float(z["params_init_sumsq"])


def test_indices_bit_exact(case):
    z, g, x = case["z"], case["g"], case["x"]
    for l in range(g.n_levels):
        idx, _ = O.grid_corners(g, x, l)
        assert np.array_equal(idx, z["indices"][l][:, : idx.shape[1]]), f"level {l}"  # Nearest: corner 0 (the cell itself) only


def test_encoded_features_bit_exact(case):
    z, g, m = case["z"], case["g"], case["m"]
    enc = O.grid_forward(g, case["x"], case["p"].astype(np.float16)[m.n_params:])
    ref = z["encoded_soa"].T[:, : g.n_features]
    assert np.array_equal(enc.view(np.uint16), ref.view(np.uint16))


def test_mlp_loss_gradients_within_tolerance(case):
    z, g, m, n_out, B = case["z"], case["g"], case["m"], case["n_out"], case["B"]
    p16 = case["p"].astype(np.float16)
    enc = np.zeros((B, m.n_input), np.float16)
    enc[:, : g.n_features] = O.grid_forward(g, case["x"], p16[m.n_params:])
    y, hidden = O.mlp_forward(m, p16[: m.n_params], enc)
    # the oracle accumulates like the reference (fp16, rounded every 16-deep step): a handful of last-place differences remain
    assert common.rel_l2(y, z["output_padded"]) < 1e-3
    assert np.count_nonzero(y.astype(np.float32) != z["output_padded"].astype(np.float32)) <= 2e-3 * y.size
    assert common.rel_l2(y[:, :n_out].astype(np.float32), z["output"]) < 1e-3
    y32, _ = O.mlp_forward(m, p16[: m.n_params], enc, acc=np.float32)  # exact accumulation: the north-star tolerance
    assert common.rel_l2(y32[:, :n_out].astype(np.float32), z["output"]) < (1.5e-2 if g.max_level < 1 else 1e-2)
    relative = case["cfg"]["loss"]["otype"] == "RelativeL2"
    # loss + gradient restatement checked on the reference's own prediction: must agree to fp32 rounding
    lv, dy = O.loss_l2(z["output_padded"], case["t"], n_out, relative=relative)
    assert abs(lv - float(z["loss"])) < 1e-5 * abs(lv)
    assert common.rel_l2(dy, z["dL_doutput"]) < 1e-6
    dW, dx = O.mlp_backward(m, p16[: m.n_params], enc, hidden, O.loss_l2(y, case["t"], n_out, relative=relative)[1])
    assert common.rel_l2(dW, z["grads_mlp"]) < 1e-2
    gg = O.grid_backward(g, case["x"], dx[:, : g.n_features])
    nz = z["grads_grid_nz_idx"].astype(np.int64)
    assert common.rel_l2(gg[nz], z["grads_grid_nz_val"]) < 1e-2  # north-star bar; 5-9e-4 measured (no ReLU mask differs any more)
    assert abs(np.count_nonzero(gg) - len(nz)) <= 2e-3 * len(nz) + 4
    if case["name"] in common.ACCUMULATE_CASES:
        # a second backward pass in GradientMode::Accumulate adds the same gradients once more (object.h:115-119; fully_fused_mlp.cu:769
        # beta = 1, grid.h:857-859 no memset): twice the first pass up to the fp16 rounding of the sum and the order of the atomics
        assert common.rel_l2(z["grads_accumulated_mlp"].astype(np.float64), 2 * z["grads_mlp"].astype(np.float64)) < 2e-3
        assert common.rel_l2(z["grads_accumulated_grid_nz_val"].astype(np.float64), 2 * z["grads_grid_nz_val"].astype(np.float64)) < 2e-3


def test_adam_first_step(case):
    z, m = case["z"], case["m"]
    o = case["cfg"]["optimizer"]
    ac = O.AdamConfig(o.get("learning_rate", 1e-3), o.get("beta1", 0.9), o.get("beta2", 0.999), o.get("epsilon", 1e-8), o.get("l2_reg", 1e-8))
    n = m.n_params
    master = case["p"][:n].copy()
    mm, vv, ss = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.uint32)
    # the reference's own gradients -> its own update (the Accumulate dump stepped on the gradients of both passes)
    O.adam_step(ac, n, master, z["grads_accumulated_mlp"] if case["name"] in common.ACCUMULATE_CASES else z["grads_mlp"], mm, vv, ss)
    assert np.max(np.abs(master - z["params_after_mlp"])) < 2e-7 + 1e-5 * ac.learning_rate


@pytest.mark.parametrize("name", ["curve_cfg1_2d_T19", "curve_nerf_3d_T19"])
def test_batch_generator_matches_reference_run(name):
    """random_uniform / synth_target restate generate_random_uniform (random.h:39-71) and the driver's target kernel:
    the first and the 1000th batch of the reference's loss-curve run are reproduced bit for bit (targets to 1 ulp of sinf)."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    B, n_in, n_out, steps, seed = (int(g[k]) for k in ("B", "n_in", "n_out", "steps", "seed"))
    rng = O.Pcg32(seed)
    x0 = O.random_uniform(rng, B * n_in).reshape(B, n_in)
    rng.advance(B * n_in * (steps - 2))
    xl = O.random_uniform(rng, B * n_in).reshape(B, n_in)
    k = len(g["x_first"])
    assert np.array_equal(x0[:k], g["x_first"]) and np.array_equal(xl[-k:], g["x_last"])
    assert np.abs(O.synth_target(x0[:k], n_out) - g["t_first"]).max() < 1e-5
    assert np.abs(O.synth_target(xl[-k:], n_out) - g["t_last"]).max() < 1e-5
    losses = g["losses"]
    assert losses.shape == (steps,) and losses[-50:].mean() < 1e-3 * losses[0]


def test_grid_input_gradient_restatement_matches_finite_differences():
    """oracle.grid_backward_input (grid.h:171-211 + 322-349) against central differences of a float64 interpolation of the
    same fp16 table; samples whose +-eps probes cross a cell boundary are excluded."""
    g = O.GridConfig(3, 5, 2, 12, 4, 1.7)
    rng = np.random.default_rng(0)
    table = (rng.standard_normal(g.n_params) * 0.3).astype(np.float16)
    x = (rng.random((256, 3)) * 0.9 + 0.05).astype(np.float32)
    dy = rng.standard_normal((256, 10)).astype(np.float16)
    dx = O.grid_backward_input(g, x, table, dy)

    def f(xx):
        tot = np.zeros(len(xx))
        t = table.astype(np.float64).reshape(-1, 2)
        for l in range(5):
            idx, _ = O.grid_corners(g, xx.astype(np.float32), l)
            p = np.float64(g.scales[l]) * xx + 0.5
            fr = p - np.floor(p)
            val = np.zeros((len(xx), 2))
            for c in range(8):
                wc = np.ones(len(xx))
                for d in range(3):
                    wc *= fr[:, d] if (c >> d) & 1 else 1 - fr[:, d]
                val += wc[:, None] * t[int(g.offsets[l]) + idx[:, c].astype(np.int64)]
            tot += (val * dy[:, 2 * l:2 * l + 2].astype(np.float64)).sum(1)
        return tot

    eps, num = 1e-4, np.zeros((256, 3))
    for d in range(3):
        xp, xm = x.astype(np.float64).copy(), x.astype(np.float64).copy()
        xp[:, d] += eps; xm[:, d] -= eps
        num[:, d] = (f(xp) - f(xm)) / (2 * eps)
    ok = (np.abs(num - dx) < 1e-2 * np.abs(num).max()).all(1)
    assert ok.mean() > 0.97
    assert np.linalg.norm((num - dx)[ok]) / np.linalg.norm(num[ok]) < 1e-2


# ---- 128-neuron networks and the OneBlob encoding (BASELINE configs #2, #4, #5) -----------------------------------
WIDE_NAMES = ["wide_dp_3d_T15", "wide_f4_3d_T14", "oneblob_2d_128x5", "oneblob16_2d_64x4", "oneblob_2d_64x5",
              "act_a_sigmoid_exp_64", "act_b_tanh_sigmoid_128", "act_c_squareplus_softplus_64", "act_d_leaky_leaky_128"]


@pytest.fixture(scope="module", params=WIDE_NAMES)
def wide_case(request):
    z = np.load(os.path.join(GOLDEN, request.param + ".npz"))
    cfg = json.loads(str(z["config"]))
    n_in, n_out, B = int(z["n_in"]), int(z["n_out"]), int(z["B"])
    g, m = common.oracle_configs(cfg, n_in, n_out, scales=z["level_scale"] if "level_scale" in z.files else None)
    x, t = common.synth_inputs(B, n_in, n_out, seed=int(z["seed"]))
    p = O.init_params(m, g, 1337, True)
    return dict(z=z, cfg=cfg, g=g, m=m, x=x, t=t, p=p, n_in=n_in, n_out=n_out, B=B)


def test_wide_param_init_bit_exact(wide_case):
    z, p, m = wide_case["z"], wide_case["p"], wide_case["m"]
    assert len(p) == int(z["n_params"])
    assert np.array_equal(p[: m.n_params], z["params_init_mlp"])
    assert np.array_equal(p[::997], z["params_init_sample"])


def test_wide_encoding_matches_reference(wide_case):
    """Grid features bit-exact; OneBlob (incl. the reference's width-64 sub-warp shuffle at n_bins = 64, where bins 31 and 63
    take their right edge from bins 0 and 32) to one fp16 ulp of the largest value: nvcc contracts the quartic CDF's
    multiply-adds, numpy does not."""
    z, cfg, g, m = wide_case["z"], wide_case["cfg"], wide_case["g"], wide_case["m"]
    enc = common.oracle_encode(cfg, g, m, wide_case["x"], wide_case["p"].astype(np.float16)[m.n_params:])
    ref = z["encoded_soa"].T
    if common.is_oneblob(cfg):
        assert enc.shape == ref.shape
        d = np.abs(enc.astype(np.float32) - ref.astype(np.float32))
        assert d.max() <= 1e-3 * np.abs(ref.astype(np.float32)).max()
        assert np.count_nonzero(d) < 5e-3 * d.size
        assert np.all(ref[:, wide_case["n_in"] * cfg["encoding"]["n_bins"]:] == 1)
    else:
        assert np.array_equal(enc[:, : g.n_features].view(np.uint16), ref[:, : g.n_features].view(np.uint16))


def test_wide_mlp_vs_reference(wide_case):
    """With the reference's arithmetic (accumulator rounded to fp16 every 16-deep MMA step, the oracle's default and what the
    B200 kernels do in TMEM) the oracle reproduces the reference's output to 1e-3; exact (fp32) accumulation stays within
    the north-star tolerance only - the gap is the reference's own fp16 accumulation."""
    z, cfg, m, n_out = wide_case["z"], wide_case["cfg"], wide_case["m"], wide_case["n_out"]
    p16 = wide_case["p"].astype(np.float16)
    enc = np.ascontiguousarray(z["encoded_soa"].T)  # the reference's own encoding output
    y, hidden = O.mlp_forward(m, p16[: m.n_params], enc)
    assert common.rel_l2(y[:, :n_out].astype(np.float32), z["output"]) < 1e-3
    y32, _ = O.mlp_forward(m, p16[: m.n_params], enc, acc=np.float32)
    assert common.rel_l2(y32[:, :n_out].astype(np.float32), z["output"]) < 2e-2
    relative = cfg["loss"]["otype"] == "RelativeL2"
    lv, dy = O.loss_l2(z["output_padded"], wide_case["t"], n_out, relative=relative)
    assert abs(lv - float(z["loss"])) < 1e-5 * abs(lv)
    assert common.rel_l2(dy, z["dL_doutput"]) < 1e-6
    dW, _ = O.mlp_backward(m, p16[: m.n_params], enc, hidden, O.loss_l2(y, wide_case["t"], n_out, relative=relative)[1], y=y)
    assert common.rel_l2(dW, z["grads_mlp"].astype(np.float64)) < 1e-2  # north-star bar for parameter gradients
    pos = m.n_params
    for k, (r, c) in enumerate(reversed(m.shapes())):  # per matrix: the reference's own fp16 split-K accumulation is the residual
        pos -= r * c
        # input matrix behind a freshly initialised grid: features ~1e-4, so its gradient entries are fp16 subnormals (~4e-6, spacing
        # 6e-8) and the reference's fp16-accumulating split-K GEMM underflows on top of that
        bar = 1e-1 if (k == len(m.shapes()) - 1 and wide_case["g"] is not None) else 3e-2
        assert common.rel_l2(dW[pos:pos + r * c], z["grads_mlp"][pos:pos + r * c].astype(np.float64)) < bar, f"dW {r}x{c}"


def test_oneblob_backward_input_matches_finite_differences():
    rng = np.random.default_rng(5)
    x = rng.random((8, 2), dtype=np.float32)
    nb = 16
    dy = rng.standard_normal((8, 32)).astype(np.float16)
    got = O.oneblob_backward_input(x, nb, dy)

    def f(xx):
        j = np.arange(2 * nb)
        lb = ((j % nb) / nb)[None, :]
        xv = xx[:, j // nb]
        cdf = lambda d: np.clip(15 / 16 * (d * nb) * (1 - 2 / 3 * (d * nb) ** 2 + 1 / 5 * (d * nb) ** 4) + 0.5, 0, 1)  # noqa: E731
        w = lambda b: cdf(b - xv) + cdf(b - xv - 1) + cdf(b - xv + 1)  # noqa: E731
        return ((w(lb + 1 / nb) - w(lb)) * dy.astype(np.float64)).sum(1)

    for d in range(2):
        xp, xm = x.astype(np.float64).copy(), x.astype(np.float64).copy()
        xp[:, d] += 1e-5; xm[:, d] -= 1e-5
        fd = (f(xp) - f(xm)) / 2e-5
        assert np.allclose(got[:, d], fd, rtol=2e-3, atol=2e-3)
