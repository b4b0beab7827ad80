"""GPU: the stand-alone grid encoding in fp32 precision (tcnn.Encoding(dtype=torch.float32); create_encoding<float>,
cpp_api.cu:156-165 of the reference) through the C ABI and the torch module, against the oracle and against dumps of the
reference's GridEncoding<float> (tests/golden/enc32_*.npz, `ref_driver enc32`)."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from oracle import tcnn_oracle as O  # noqa: E402
from tests import common  # noqa: E402
from tests.test_oracle_encoding_fp32 import CASES, load_fp32_case  # noqa: E402


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device: the gpu-marked tests need a B200 (there is no CPU fallback)")
    return torch.device("cuda:0")


@pytest.mark.parametrize("name", sorted(CASES))
def test_c_abi_fp32_vs_reference_dump_and_oracle(dev, name):
    import tinycudann as tcnn
    n_in, enc = CASES[name]
    native = tcnn.NativeModule(n_in, 0, {"encoding": enc, "precision": "fp32"})
    g, z, (x_np, table_np, dy_np) = load_fp32_case(name, np.array(native.info.level_scale[: native.info.n_levels], dtype=np.float32))
    assert native.param_dtype == torch.float32 and native.info.encoded_width == dy_np.shape[1]
    x, table, dy = (torch.from_numpy(a).to(dev) for a in (x_np, table_np, dy_np))
    ws, y = native.fwd(x, table, need_ctx=True)
    assert y.dtype == torch.float32
    F = g.n_features
    y_np = y.cpu().numpy()
    assert common.rel_l2(y_np, z["encoded"]) < 1e-6 and not y_np[:, F:].any()
    assert np.abs(y_np[:, :F] - O.grid_forward_f32(g, x_np, table_np)).max() < 1e-6
    _, y_inf = native.fwd(x, table, need_ctx=False)
    assert torch.equal(y_inf, y)
    dx, gt = native.bwd(ws, x, table, dy, want_params_grad=True, want_input_grad=True)
    assert gt.dtype == torch.float32
    assert common.rel_l2(gt.cpu().numpy(), z["dL_dtable"]) < 1e-5  # fp32 atomics on both sides
    assert common.rel_l2(gt.cpu().numpy(), O.grid_backward_f32(g, x_np, dy_np)) < 1e-5
    if g.nearest:
        assert not dx.cpu().numpy().any() and not z["dL_dx"].any()
    else:
        assert common.rel_l2(dx.cpu().numpy(), z["dL_dx"]) < 1e-5
        assert common.rel_l2(dx.cpu().numpy(), O.grid_backward_input(g, x_np, table_np, dy_np, fp32=True)) < 1e-5


def test_encoding_module_fp32_autograd(dev):
    """tcnn.Encoding(dtype=torch.float32): fp32 parameters and output, loss scale 1, gradients against the oracle."""
    import tinycudann as tcnn
    n_in, enc = CASES["hash3d_smooth"]
    module = tcnn.Encoding(n_in, enc, dtype=torch.float32).to(dev)
    assert module.dtype == torch.float32 and module.loss_scale == 1.0 and module.params.dtype == torch.float32
    native = module.native_tcnn_module
    g, _, (x_np, table_np, dy_np) = load_fp32_case("hash3d_smooth", np.array(native.info.level_scale[: native.info.n_levels], dtype=np.float32))
    with torch.no_grad():
        module.params.copy_(torch.from_numpy(table_np).to(dev))
    B = 200  # not a multiple of 256: padded inside, sliced outside
    x = torch.from_numpy(x_np[:B]).to(dev).requires_grad_(True)
    y = module(x)
    assert y.dtype == torch.float32 and y.shape == (B, g.n_features)
    c = torch.from_numpy(dy_np[:B, : g.n_features]).to(dev)
    (y * c).sum().backward()
    assert np.abs(y.detach().cpu().numpy() - O.grid_forward_f32(g, x_np[:B], table_np)).max() < 1e-6
    assert common.rel_l2(module.params.grad.cpu().numpy(), O.grid_backward_f32(g, x_np[:B], dy_np[:B])) < 1e-5
    assert common.rel_l2(x.grad.cpu().numpy(), O.grid_backward_input(g, x_np[:B], table_np, dy_np[:B], fp32=True)) < 1e-5


def test_fp32_precision_is_for_standalone_encodings(dev):
    import tinycudann as tcnn
    with pytest.raises(RuntimeError, match="stand-alone encoding"):
        tcnn.NativeModule(2, 3, dict(common.CONFIG_1, precision="fp32"))
