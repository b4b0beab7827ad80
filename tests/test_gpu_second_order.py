"""GPU: second-order gradients of the stand-alone grid encoding (encodings/grid.h:351-650; scripts/test_grid_bwdbwd.py of the
reference) through tcnn.Encoding's double backward and through the C ABI, against the oracle restatement - which
tests/test_oracle_second_order.py checks against torch autograd of an independent float64 implementation."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from oracle import tcnn_oracle as O  # noqa: E402
from tests import common  # noqa: E402

ENCODINGS = common.SECOND_ORDER_ENCODINGS


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device: the gpu-marked tests need a B200 (there is no CPU fallback)")
    return torch.device("cuda:0")


def _oracle_cfg(native, D, enc):
    L = native.info.n_levels
    return common.oracle_configs({"encoding": enc, "network": {"n_neurons": 64, "n_hidden_layers": 1}}, D, 1,
                                 scales=np.array(native.info.level_scale[:L], dtype=np.float32))[0]


@pytest.mark.parametrize("name", sorted(ENCODINGS))
def test_c_abi_backward_backward_input(dev, name):
    import tinycudann as tcnn
    D, enc = ENCODINGS[name]
    native = tcnn.NativeModule(D, 0, {"encoding": enc})
    g = _oracle_cfg(native, D, enc)
    B, W = 512, native.info.encoded_width
    rng = np.random.default_rng(3)
    x_np = rng.random((B, D), dtype=np.float32)
    table_np = rng.uniform(-1, 1, g.n_params).astype(np.float16)
    dy_np = np.zeros((B, W), np.float16)
    dy_np[:, : g.n_features] = rng.standard_normal((B, g.n_features)).astype(np.float16)
    v_np = rng.standard_normal((B, D)).astype(np.float32)
    x, table, dy, v = (torch.from_numpy(a).to(dev) for a in (x_np, table_np, dy_np, v_np))
    dd, gp, dx = native.bwd_bwd_input(x, table, v, dy)
    ref_dd = O.grid_bwd_bwd_dLddLdy(g, x_np, table_np, v_np)
    assert common.rel_l2(dd.cpu().numpy()[:, : g.n_features], ref_dd) < 2e-3
    assert not dd.cpu().numpy()[:, g.n_features:].any()
    ref_g = O.grid_bwd_bwd_grid(g, x_np, dy_np[:, : g.n_features], v_np)
    assert common.rel_l2(gp.cpu().numpy().astype(np.float64), ref_g) < 1e-2  # fp16 atomics
    ref_dx = O.grid_bwd_bwd_input(g, x_np, table_np, dy_np[:, : g.n_features], v_np)
    if np.linalg.norm(ref_dx) > 0:
        assert common.rel_l2(dx.cpu().numpy(), ref_dx) < 1e-4
    else:  # 2-D... never: the cross terms exist for D >= 2; kept for a 1-D Linear grid where the Hessian vanishes
        assert not dx.cpu().numpy().any()


@pytest.mark.parametrize("name", sorted(ENCODINGS))
def test_encoding_double_backward_matches_oracle(dev, name):
    """tcnn.Encoding: eikonal-style loss on the spatial gradient of the encoded field."""
    import tinycudann as tcnn
    D, enc = ENCODINGS[name]
    module = tcnn.Encoding(D, enc).to(dev)
    native = module.native_tcnn_module
    g = _oracle_cfg(native, D, enc)
    B = 256
    rng = np.random.default_rng(9)
    x_np = rng.random((B, D), dtype=np.float32)
    table_np = rng.uniform(-1, 1, g.n_params).astype(np.float16)
    with torch.no_grad():
        module.params.copy_(torch.from_numpy(table_np.astype(np.float32)).to(dev))
    c_np = rng.standard_normal((B, g.n_features)).astype(np.float16)  # dL/dy of the first backward = c (loss = sum(y * c))
    v_np = rng.standard_normal((B, D)).astype(np.float32)
    x = torch.from_numpy(x_np).to(dev).requires_grad_(True)
    c = torch.from_numpy(c_np).to(dev).requires_grad_(True)
    y = module(x)
    assert y.shape == (B, g.n_features)
    (dL_dx,) = torch.autograd.grad((y * c).sum(), x, create_graph=True)
    assert common.rel_l2(dL_dx.detach().cpu().numpy(), O.grid_backward_input(g, x_np, table_np, c_np)) < 2e-3
    S = (dL_dx * torch.from_numpy(v_np).to(dev)).sum()
    g_c, g_params, g_x = torch.autograd.grad(S, (c, module.params, x), allow_unused=True)
    assert common.rel_l2(g_c.float().cpu().numpy(), O.grid_bwd_bwd_dLddLdy(g, x_np, table_np, v_np)) < 3e-3
    assert common.rel_l2(g_params.float().cpu().numpy().astype(np.float64), O.grid_bwd_bwd_grid(g, x_np, c_np, v_np)) < 2e-2
    assert common.rel_l2(g_x.cpu().numpy(), O.grid_bwd_bwd_input(g, x_np, table_np, c_np, v_np)) < 2e-3


@pytest.mark.parametrize("name", sorted(ENCODINGS))
def test_c_abi_vs_reference_dump(dev, name):
    """The same seeded inputs the reference's GridEncoding was run on (tests/golden/bwdbwd_*.npz, `ref_driver bwdbwd`)."""
    import tinycudann as tcnn
    from tests.test_oracle_second_order import load_second_order_case
    D, enc = ENCODINGS[name]
    g, z, (x_np, table_np, dy_np, v_np) = load_second_order_case(name)
    native = tcnn.NativeModule(D, 0, {"encoding": enc})
    W = native.info.encoded_width
    dy_pad = np.zeros((x_np.shape[0], W), np.float16)
    dy_pad[:, : g.n_features] = dy_np[:, : g.n_features]
    x, table, dy, v = (torch.from_numpy(a).to(dev) for a in (x_np, table_np, dy_pad, v_np))
    dd, gp, dx = native.bwd_bwd_input(x, table, v, dy)
    F = g.n_features
    assert common.rel_l2(dd.cpu().numpy()[:, :F], z["dL2_ddLdy"][:, :F]) < 1e-3
    assert common.rel_l2(gp.cpu().numpy(), z["dL2_dtable"]) < 5e-3  # fp16 atomics on both sides
    assert common.rel_l2(dx.cpu().numpy(), z["dL2_dx"]) < 1e-5
