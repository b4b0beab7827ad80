"""CPU: the oracle's fp32-precision grid encoding (kernel_grid / kernel_grid_backward with T = float) against dumps of the
reference's GridEncoding<float> (`ref_driver enc32`, tests/golden/enc32_*.npz)."""
import os

import numpy as np
import pytest

from oracle import tcnn_oracle as O
from tests import common

CASES = common.FP32_ENCODINGS

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_fp32_case(name, scales=None):
    """scales: the device-evaluated level scales (GPU tests pass them; a host exp2 can be one ulp off CUDA's exp2f, which moves an
    fp32 feature by up to ~1e-5 - invisible after fp16 rounding, visible here)."""
    n_in, enc = CASES[name]
    z = np.load(os.path.join(GOLDEN, f"enc32_{name}.npz"))
    g = common.oracle_configs({"encoding": enc, "network": {"n_neurons": 64, "n_hidden_layers": 1}}, n_in, 1, scales=scales)[0]
    W = (g.n_features + 7) // 8 * 8
    assert z["encoded"].shape == (int(z["B"]), W) and z["dL_dtable"].size == g.n_params
    return g, z, common.fp32_inputs(n_in, g.n_params, g.n_features, W, int(z["B"]), int(z["seed"]))


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_fp32_encoding_vs_reference_dump(name):
    g, z, (x, table, dy) = load_fp32_case(name)
    F = g.n_features
    y = O.grid_forward_f32(g, x, table)
    assert np.abs(y - z["encoded"][:, :F]).max() < 3e-5 and not z["encoded"][:, F:].any()  # host-evaluated level scales, see above
    assert common.rel_l2(O.grid_backward_f32(g, x, dy), z["dL_dtable"]) < 1e-5
    dx = O.grid_backward_input(g, x, table, dy, fp32=True)
    if g.nearest:
        assert not dx.any() and not z["dL_dx"].any()
    else:
        assert common.rel_l2(dx, z["dL_dx"]) < 1e-5
