"""CPU: the oracle's restatement of the second-order grid kernels (encodings/grid.h:351-650) against torch autograd through an
independent float64 implementation of the grid encoding - the check scripts/test_grid_bwdbwd.py of the reference performs."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
from oracle import tcnn_oracle as O  # noqa: E402


def torch_grid_encode(cfg, x, table):
    """float64, differentiable w.r.t. x and table; indices from the oracle's integer index function."""
    B, D = x.shape
    F = cfg.n_features_per_level
    t = table.reshape(-1, F)
    outs = []
    for l in range(cfg.n_levels):
        scale = float(cfg.scales[l])
        pos = x * scale + 0.5
        cell = torch.floor(pos).detach()
        frac = pos - cell
        w = frac * frac * (3 - 2 * frac) if cfg.smooth else frac
        cell_np = (cell.numpy().astype(np.int64) & 0xFFFFFFFF).astype(np.uint32)
        res = torch.zeros(B, F, dtype=torch.float64)
        for c in range(1 << D):
            weight = torch.ones(B, dtype=torch.float64)
            pg = np.empty((B, D), np.uint32)
            for d in range(D):
                if (c >> d) & 1:
                    weight = weight * w[:, d]; pg[:, d] = cell_np[:, d] + np.uint32(1)
                else:
                    weight = weight * (1 - w[:, d]); pg[:, d] = cell_np[:, d]
            idx = O.grid_index(pg, cfg.level_size(l), int(cfg.resolutions[l]), cfg.hashed).astype(np.int64) + int(cfg.offsets[l])
            res = res + weight[:, None] * t[torch.from_numpy(idx)]
        outs.append(res)
    return torch.cat(outs, dim=1)


@pytest.mark.parametrize("D,interp,gtype", [(2, "Linear", "Hash"), (3, "Linear", "Hash"), (2, "Smoothstep", "Hash"), (3, "Smoothstep", "Tiled")])
def test_second_order_restatement_matches_autograd(D, interp, gtype):
    rng = np.random.default_rng(5)
    cfg = O.GridConfig(D, n_levels=4, n_features_per_level=2, log2_hashmap_size=8, base_resolution=4, per_level_scale=1.7, grid_type=gtype, interpolation=interp)
    B = 64
    x_np = rng.random((B, D), dtype=np.float32)
    table_np = rng.uniform(-1, 1, cfg.n_params).astype(np.float16)
    dy_np = rng.standard_normal((B, cfg.n_features)).astype(np.float16)
    v_np = rng.standard_normal((B, D)).astype(np.float32)

    x = torch.tensor(x_np.astype(np.float64), requires_grad=True)
    table = torch.tensor(table_np.astype(np.float64), requires_grad=True)
    dy = torch.tensor(dy_np.astype(np.float64), requires_grad=True)
    y = torch_grid_encode(cfg, x, table)
    (dL_dx,) = torch.autograd.grad((y * dy).sum(), x, create_graph=True)
    assert np.allclose(dL_dx.detach().numpy(), O.grid_backward_input(cfg, x_np, table_np, dy_np), rtol=2e-3, atol=2e-3)
    S = (dL_dx * torch.tensor(v_np.astype(np.float64))).sum()
    g_dy, g_table, g_x = torch.autograd.grad(S, (dy, table, x), allow_unused=True)

    got_dy = O.grid_bwd_bwd_dLddLdy(cfg, x_np, table_np, v_np).astype(np.float64)
    assert np.linalg.norm(got_dy - g_dy.numpy()) <= 2e-3 * np.linalg.norm(g_dy.numpy())
    got_table = O.grid_bwd_bwd_grid(cfg, x_np, dy_np, v_np)
    assert np.linalg.norm(got_table - g_table.numpy()) <= 3e-3 * np.linalg.norm(g_table.numpy())  # fp16 weights and products
    got_x = O.grid_bwd_bwd_input(cfg, x_np, table_np, dy_np, v_np).astype(np.float64)
    ref_x = np.zeros_like(got_x) if g_x is None else g_x.numpy()
    assert np.linalg.norm(got_x - ref_x) <= 2e-4 * max(np.linalg.norm(ref_x), 1.0)


# --- pinned to the reference: dumps of its own GridEncoding::backward_backward_input (ref_driver bwdbwd, make_golden.py) ---
import os  # noqa: E402

from tests import common  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_second_order_case(name):
    n_in, enc = common.SECOND_ORDER_ENCODINGS[name]
    z = np.load(os.path.join(GOLDEN, f"bwdbwd_{name}.npz"))
    g = common.oracle_configs({"encoding": enc, "network": {"n_neurons": 64, "n_hidden_layers": 1}}, n_in, 1)[0]
    W = (g.n_features + 7) // 8 * 8
    assert z["dL2_ddLdy"].shape == (256, W) and z["dL2_dtable"].size == g.n_params
    return g, z, common.second_order_inputs(n_in, g.n_params, g.n_features, W, int(z["B"]), int(z["seed"]))


@pytest.mark.parametrize("name", sorted(common.SECOND_ORDER_ENCODINGS))
def test_oracle_second_order_vs_reference_dump(name):
    g, z, (x, table, dy, v) = load_second_order_case(name)
    F = g.n_features
    assert common.rel_l2(O.grid_backward_input(g, x, table, dy[:, :F]), z["dL_dx"]) < 1e-5
    assert common.rel_l2(O.grid_bwd_bwd_dLddLdy(g, x, table, v), z["dL2_ddLdy"][:, :F]) < 5e-5  # fp16 result
    assert not z["dL2_ddLdy"][:, F:].any()
    assert common.rel_l2(O.grid_bwd_bwd_grid(g, x, dy[:, :F], v).reshape(-1), z["dL2_dtable"]) < 3e-3  # fp16 atomics in the reference
    assert common.rel_l2(O.grid_bwd_bwd_input(g, x, table, dy[:, :F], v), z["dL2_dx"]) < 1e-5
