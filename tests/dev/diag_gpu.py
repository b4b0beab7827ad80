"""Bring-up diagnostic (run on the GPU box, not a pytest): each kernel against the oracle, verbose numbers.
usage: python tests/dev/diag_gpu.py [config_name] [B]"""
import os
import sys
import time
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tiny-cuda-nn_b200"))

import tinycudann as tcnn  # noqa: E402
from tinycudann import _lib  # noqa: E402
from tinycudann._lib import check, ptr  # noqa: E402
from oracle import tcnn_oracle as O  # noqa: E402
from tests import common  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "CONFIG_1"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
cfg = getattr(common, name)
n_in = 3 if name == "CONFIG_NERF" or name == "CONFIG_F4" else 2
n_out = 3
dev = torch.device("cuda:0")
S = lambda: _lib.C.c_void_p(torch.cuda.current_stream().cuda_stream)


def step(title, fn):
    t0 = time.time()
    try:
        r = fn()
        torch.cuda.synchronize()
        print(f"[ok]   {title}: {r}  ({time.time() - t0:.2f}s)", flush=True)
        return True
    except Exception as e:  # noqa: BLE001
        print(f"[FAIL] {title}: {e}", flush=True)
        traceback.print_exc()
        return False


native = tcnn.NativeModule(n_in, n_out, cfg)
info = native.info
L = info.n_levels
scales = np.array(info.level_scale[:L], dtype=np.float32)
g, m = common.oracle_configs(cfg, n_in, n_out, scales=scales)
print("n_params", info.n_params, "mlp", info.n_params_mlp, "grid", info.n_params_grid, "E", info.encoded_width, flush=True)
print("scales dev", scales, "\nscales host-numpy", common.oracle_configs(cfg, n_in, n_out)[0].scales, flush=True)
assert g.n_params == info.n_params_grid and m.n_params == info.n_params_mlp

x_np, t_np = common.synth_inputs(B, n_in, n_out, seed=1)
x = torch.from_numpy(x_np).to(dev)
t = torch.from_numpy(t_np).to(dev)
lib = native.lib
state = {}


def t_init():
    p = native.initial_params(1337, dev, trainer_seeding=True)
    ref = O.init_params(m, g, 1337, True)
    got = p.cpu().numpy()
    state["p32"] = got
    return f"bit-exact={np.array_equal(got, ref)} maxdiff={np.abs(got - ref).max():.3e}"


step("init_params vs oracle", t_init)
# make the grid values larger so that features are well conditioned for the comparisons
rng = np.random.default_rng(7)
p32 = state["p32"].copy()
p32[m.n_params:] = rng.uniform(-0.5, 0.5, g.n_params).astype(np.float32)
p16_np = p32.astype(np.float16)
p16 = torch.from_numpy(p16_np).to(dev)


def t_indices():
    out = torch.empty(L * B * (1 << n_in), dtype=torch.int32, device=dev)
    check(lib.tcnn_b200_grid_indices(native.handle, S(), B, ptr(x), ptr(out)))
    got = out.cpu().numpy().view(np.uint32).reshape(L, B, 1 << n_in)
    bad = 0
    for l in range(L):
        idx, _ = O.grid_corners(g, x_np, l)
        bad += int((idx != got[l]).sum())
    return f"mismatching indices={bad} of {got.size}"


step("grid indices vs oracle", t_indices)


def t_encode():
    enc = torch.empty(B, info.encoded_width, dtype=torch.half, device=dev)
    check(lib.tcnn_b200_encode_forward(native.handle, S(), B, ptr(x), ptr(p16[m.n_params:]), ptr(enc)))
    got = enc.cpu().numpy()
    ref = O.grid_forward(g, x_np, p16_np[m.n_params:])
    state["enc"] = got
    nf = g.n_features
    return f"bit-exact={np.array_equal(got[:, :nf].view(np.uint16), ref.view(np.uint16))} rel_l2={common.rel_l2(got[:, :nf], ref):.3e} pad_zero={not got[:, nf:].any()}"


step("encode forward vs oracle", t_encode)


def t_mlp_fwd():
    enc = torch.from_numpy(state["enc"]).to(dev)
    out = torch.empty(B, 16, dtype=torch.half, device=dev)
    check(lib.tcnn_b200_mlp_forward(native.handle, S(), B, ptr(enc), ptr(p16), ptr(out)))
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    y, hidden = O.mlp_forward(m, p16_np[: m.n_params], state["enc"])
    state["y"], state["hidden"] = y, hidden
    return f"rel_l2={common.rel_l2(got, y):.3e} max|y|={np.abs(y).max():.3f} got[0,:4]={got[0, :4]} ref[0,:4]={y[0, :4]}"


step("mlp forward vs oracle", t_mlp_fwd)


def t_mlp_bwd():
    enc = torch.from_numpy(state["enc"]).to(dev)
    dy_np = (rng.standard_normal((B, 16)) * 0.1).astype(np.float16)
    dy_np[:, n_out:] = 0
    dy = torch.from_numpy(dy_np).to(dev)
    ws = native.workspace(B, dev)
    gp = torch.empty(m.n_params, dtype=torch.half, device=dev)
    dx = torch.empty(B, info.encoded_width, dtype=torch.half, device=dev)
    check(lib.tcnn_b200_mlp_backward(native.handle, S(), B, ptr(enc), ptr(p16), ptr(dy), ptr(gp), ptr(dx), ptr(ws)))
    torch.cuda.synchronize()
    dparams, dx_ref = O.mlp_backward(m, p16_np[: m.n_params], state["enc"], state["hidden"], dy_np)
    got_g = gp.cpu().numpy().astype(np.float64)
    res = [f"dx rel_l2={common.rel_l2(dx.cpu().numpy(), dx_ref):.3e}"]
    pos = 0
    for (r, c) in m.shapes():
        res.append(f"dW[{r}x{c}] rel_l2={common.rel_l2(got_g[pos:pos + r * c], dparams[pos:pos + r * c]):.3e}")
        pos += r * c
    return " ".join(res)


step("mlp backward vs oracle", t_mlp_bwd)


def t_train():
    ws = native.workspace(B, dev)
    grads = torch.zeros(info.n_params, dtype=torch.half, device=dev)
    loss = torch.zeros(1, dtype=torch.float32, device=dev)
    out = torch.empty(B, 16, dtype=torch.half, device=dev)
    check(lib.tcnn_b200_training_fwd_bwd(native.handle, S(), B, ptr(x), ptr(t), ptr(p16), ptr(grads), 128.0, 1, ptr(loss), ptr(out), ptr(ws)))
    torch.cuda.synchronize()
    enc = O.grid_forward(g, x_np, p16_np[m.n_params:])
    encp = np.zeros((B, m.n_input), np.float16); encp[:, : g.n_features] = enc
    y, hidden = O.mlp_forward(m, p16_np[: m.n_params], encp)
    lv, dy = O.loss_l2(y, t_np, n_out, relative=(info.loss_type == 1))
    dparams, dx = O.mlp_backward(m, p16_np[: m.n_params], encp, hidden, dy)
    ggrid = O.grid_backward(g, x_np, dx[:, : g.n_features])
    got = grads.cpu().numpy().astype(np.float64)
    state["grads"] = grads
    return (f"loss got={loss.item():.6f} ref={lv:.6f} out rel_l2={common.rel_l2(out.cpu().numpy(), y):.3e} "
            f"dW rel_l2={common.rel_l2(got[: m.n_params], dparams):.3e} dGrid rel_l2={common.rel_l2(got[m.n_params:], ggrid):.3e} "
            f"nonzero grid grads got={np.count_nonzero(got[m.n_params:])} ref={np.count_nonzero(ggrid)}")


step("training fwd+bwd vs oracle", t_train)


def t_adam():
    P = info.n_params
    master = torch.from_numpy(p32).to(dev)
    params = p16.clone()
    grads = state["grads"].clone()
    m1 = torch.zeros(P, device=dev); v1 = torch.zeros(P, device=dev); st = torch.zeros(P, dtype=torch.int32, device=dev)
    g_np = grads.cpu().numpy()
    check(lib.tcnn_b200_adam_step(native.handle, S(), 128.0, 1, ptr(master), ptr(params), ptr(grads), ptr(m1), ptr(v1), ptr(st), 1))
    torch.cuda.synchronize()
    o = cfg["optimizer"]
    ac = O.AdamConfig(o.get("learning_rate", 1e-3), o.get("beta1", 0.9), o.get("beta2", 0.999), o.get("epsilon", 1e-8), o.get("l2_reg", 1e-8))
    mm = p32.copy(); m0 = np.zeros(P, np.float32); v0 = np.zeros(P, np.float32); s0 = np.zeros(P, np.uint32)
    O.adam_step(ac, m.n_params, mm, g_np, m0, v0, s0)
    got = master.cpu().numpy()
    return (f"master maxdiff={np.abs(got - mm).max():.3e} rel_l2(update)={common.rel_l2(got - p32, mm - p32):.3e} steps_equal={np.array_equal(st.cpu().numpy().view(np.uint32), s0)} "
            f"grid grads zeroed={not grads[m.n_params:].any().item()}")


step("adam step vs oracle", t_adam)
print("done", flush=True)
