"""Development diagnostic (GPU box): the 128-neuron kernels and the OneBlob encoding through the C ABI against the numpy
oracle, stage by stage, printing relative errors instead of asserting.   python tests/dev/diag_wide.py"""
import os
import sys
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tiny-cuda-nn_b200"))
from oracle import tcnn_oracle as O  # noqa: E402
from tests import common  # noqa: E402
import tinycudann as tcnn  # noqa: E402
from tinycudann._lib import C, check, ptr  # noqa: E402

dev = torch.device("cuda:0")


def S():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def per_matrix(m, got, ref):
    pos, out = 0, []
    for (r, c) in m.shapes():
        out.append(f"{r}x{c}:{common.rel_l2(got[pos:pos + r * c], ref[pos:pos + r * c]):.2e}")
        pos += r * c
    return " ".join(out)


def run(name, cfg, n_in, n_out, B):
    print(f"=== {name} B={B}", flush=True)
    native = tcnn.NativeModule(n_in, n_out, cfg)
    info, lib = native.info, native.lib
    scales = np.array(info.level_scale[: info.n_levels], dtype=np.float32)
    g, m = common.oracle_configs(cfg, n_in, n_out, scales=scales if info.n_levels else None)
    assert m.n_params == info.n_params_mlp, (m.n_params, info.n_params_mlp)
    rng = np.random.default_rng(7)
    x_np, t_np = common.synth_inputs(B, n_in, n_out, seed=3)
    p32 = O.init_params(m, g, 1337, True)
    if g is not None:  # larger table values so that the encoded features are not ~1e-4
        p32[m.n_params:] = rng.uniform(-0.5, 0.5, p32.size - m.n_params).astype(np.float32)
    p16_np = p32.astype(np.float16)
    x, t, p16 = (torch.from_numpy(a).to(dev) for a in (x_np, t_np, p16_np))
    ws = native.workspace(B, dev)
    E = info.encoded_width

    # encoding alone (through inference's first stage is not exposed for OneBlob: compare via forward below)
    enc_np = common.oracle_encode(cfg, g, m, x_np, p16_np[m.n_params:])
    enc = torch.from_numpy(enc_np).to(dev)

    # MLP forward
    out = torch.zeros(B, 16, dtype=torch.half, device=dev)
    check(lib.tcnn_b200_mlp_forward(native.handle, S(), B, ptr(enc), ptr(p16), ptr(out)))
    torch.cuda.synchronize()
    y, hidden = O.mlp_forward(m, p16_np[: m.n_params], enc_np)
    print("mlp_forward rel", common.rel_l2(out.cpu().numpy(), y), "max|y|", float(np.abs(y.astype(np.float32)).max()), flush=True)

    # MLP backward
    dy_np = (rng.standard_normal((B, 16)) * 0.1).astype(np.float16)
    dy_np[:, n_out:] = 0
    dy = torch.from_numpy(dy_np).to(dev)
    gp = torch.zeros(m.n_params, dtype=torch.half, device=dev)
    dx = torch.zeros(B, E, dtype=torch.half, device=dev)
    check(lib.tcnn_b200_mlp_backward(native.handle, S(), B, ptr(enc), ptr(p16), ptr(dy), ptr(gp), ptr(dx), ptr(ws)))
    torch.cuda.synchronize()
    dW, dx_ref = O.mlp_backward(m, p16_np[: m.n_params], enc_np, hidden, dy_np, y=y)
    print("mlp_backward dx rel", common.rel_l2(dx.cpu().numpy(), dx_ref), flush=True)
    print("mlp_backward dW ", per_matrix(m, gp.cpu().numpy().astype(np.float64), dW), flush=True)

    # full inference (encoding kernel + MLP)
    o32 = torch.zeros(B, n_out, dtype=torch.float32, device=dev)
    check(lib.tcnn_b200_inference(native.handle, S(), B, ptr(x), ptr(p16), None, ptr(o32), ptr(ws)))
    torch.cuda.synchronize()
    print("inference rel", common.rel_l2(o32.cpu().numpy(), y[:, :n_out].astype(np.float32)), flush=True)

    # training forward/backward
    grads = torch.zeros(info.n_params, dtype=torch.half, device=dev)
    loss = torch.zeros(1, dtype=torch.float32, device=dev)
    check(lib.tcnn_b200_training_fwd_bwd(native.handle, S(), B, ptr(x), ptr(t), ptr(p16), ptr(grads), 128.0, 1, ptr(loss), ptr(out), ptr(ws)))
    torch.cuda.synchronize()
    lv, dyl = O.loss_l2(y, t_np, n_out, relative=(info.loss_type == 1))
    dW, dxl = O.mlp_backward(m, p16_np[: m.n_params], enc_np, hidden, dyl, y=y)
    got = grads.cpu().numpy().astype(np.float64)
    print("train loss", loss.item(), "oracle", lv, flush=True)
    print("train dW ", per_matrix(m, got[: m.n_params], dW), flush=True)
    if g is not None:
        gg = O.grid_backward(g, x_np, dxl[:, : g.n_features])
        print("train grid grads rel", common.rel_l2(got[m.n_params:], gg.reshape(-1)), flush=True)

    # a few optimizer steps through the Trainer: loss must fall
    model = tcnn.create_from_config(n_in, n_out, cfg, device=dev)
    losses = []
    for i in range(60):
        xb, tb = common.synth_inputs(B, n_in, n_out, seed=100 + i)
        model.trainer.training_step(torch.from_numpy(xb).to(dev), torch.from_numpy(tb).to(dev))
        losses.append(model.trainer.loss())
    print("trainer losses", losses[0], losses[1], losses[-1], flush=True)


CASES = [
    ("oneblob_64x5", common.CONFIG_ONEBLOB_64, 2, 3, 1024),
    ("oneblob_64x5_big", common.CONFIG_ONEBLOB_64, 2, 3, 1 << 16),
    ("act_a", common.CONFIG_ACT_A, 2, 3, 512),
    ("act_b", common.CONFIG_ACT_B, 2, 3, 512),
    ("act_c", common.CONFIG_ACT_C, 2, 3, 512),
    ("act_d", common.CONFIG_ACT_D, 3, 3, 512),
    ("wide_dp_3d", common.CONFIG_WIDE_DP, 3, 3, 512),
    ("wide_f4_3d", common.CONFIG_WIDE_F4, 3, 4, 1024),
    ("oneblob_128x5", common.CONFIG_ONEBLOB, 2, 3, 1024),
    ("oneblob16_64x4", common.CONFIG_ONEBLOB16, 2, 3, 512),
    ("wide_dp_3d_big", common.CONFIG_WIDE_DP, 3, 3, 1 << 16),
]
if __name__ == "__main__":
    only = sys.argv[1:] or None
    for c in CASES:
        if only and c[0] not in only:
            continue
        try:
            run(*c)
        except Exception:  # noqa: BLE001
            traceback.print_exc()
            torch.cuda.synchronize()
