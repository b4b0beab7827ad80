"""Development aid: small loss curve (B=4096, 300 steps) from the reference build, ours, and (CPU) the numpy oracle."""
import json, os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tiny-cuda-nn_b200")]
from tests import common
from oracle import tcnn_oracle as O
cfg, n_in, n_out, B, steps, seed = common.CONFIG_1, 2, 3, 4096, 300, 42
out = os.path.join(ROOT, "gpurun_out", "curve_small"); os.makedirs(out, exist_ok=True)
mode = sys.argv[1]
if mode == "gpu":
    import torch, tinycudann as tcnn
    REF = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    for rep in range(2):
        with tempfile.TemporaryDirectory() as d:
            json.dump(cfg, open(os.path.join(d, "cfg.json"), "w"))
            subprocess.run([REF, "curve", os.path.join(d, "cfg.json"), str(n_in), str(n_out), str(B), str(steps), str(seed), d], capture_output=True, timeout=600)
            np.save(os.path.join(out, f"ref{rep}.npy"), np.fromfile(os.path.join(d, "losses.f32"), dtype=np.float32))
    dev = torch.device("cuda:0")
    for rep in range(2):
        rng = O.Pcg32(seed)
        tr = tcnn.create_from_config(n_in, n_out, cfg, seed=1337, device=dev).trainer
        losses = torch.zeros(steps, device=dev)
        for i in range(steps):
            x_np = O.random_uniform(rng, B * n_in).reshape(B, n_in)
            tr.training_step(torch.from_numpy(x_np).to(dev), torch.from_numpy(O.synth_target(x_np, n_out)).to(dev)); losses[i] = tr.loss_value[0]
        np.save(os.path.join(out, f"ours{rep}.npy"), losses.cpu().numpy())
else:
    acc = "hmma16" if mode.startswith("oracle16") else np.float32
    steps = 220
    g, m = common.oracle_configs(cfg, n_in, n_out)
    p32 = O.init_params(m, g, 1337, True)
    ad = O.AdamConfig()
    ad.learning_rate = 1e-3
    mm = np.zeros_like(p32); vv = np.zeros_like(p32); st = np.zeros(p32.shape, np.uint32)
    rng = O.Pcg32(seed); losses = []
    for i in range(steps):
        x = O.random_uniform(rng, B * n_in).reshape(B, n_in); t = O.synth_target(x, n_out)
        p16 = p32.astype(np.float16)
        enc = O.grid_forward(g, x, p16[m.n_params:])
        y, hidden = O.mlp_forward(m, p16[:m.n_params], enc, acc)
        lv, dy = O.loss_l2(y, t, n_out)
        dW, dx = O.mlp_backward(m, p16[:m.n_params], enc, hidden, dy, acc, "hmma16" if "dw16" in mode else np.float64)
        gg = O.grid_backward(g, x, dx, sequential_fp16="seq16" in mode)
        grads = np.concatenate([dW, gg]).astype(np.float16)
        O.adam_step(ad, m.n_params, p32, grads, mm, vv, st)
        losses.append(lv)
        if i % 50 == 0: print(i, lv, flush=True)
    np.save(os.path.join(out, f"{mode}.npy"), np.array(losses, np.float32))
