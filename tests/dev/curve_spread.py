"""Development aid: run-to-run spread of the 1000-step loss (reference build and ours), same batches."""
import json, os, subprocess, sys, tempfile
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tiny-cuda-nn_b200")]
from tests import common
from oracle import tcnn_oracle as O
import tinycudann as tcnn
REF = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
name, cfg, n_in, n_out, B, steps, seed = "cfg1", common.CONFIG_1, 2, 3, 1 << 16, 1000, 42
if len(sys.argv) > 1 and sys.argv[1] == "nerf":
    name, cfg, n_in, n_out, seed = "nerf", common.CONFIG_NERF, 3, 4, 43
wins = [(0, 10), (50, 100), (200, 500), (500, 950), (950, 1000)]
def show(tag, l): print(tag, " ".join(f"{l[a:b].mean():.4e}" for a, b in wins), flush=True)
for rep in range(4):
    with tempfile.TemporaryDirectory() as d:
        json.dump(cfg, open(os.path.join(d, "cfg.json"), "w"))
        subprocess.run([REF, "curve", os.path.join(d, "cfg.json"), str(n_in), str(n_out), str(B), str(steps), str(seed), d], capture_output=True, timeout=600)
        show(f"ref{rep} ", np.fromfile(os.path.join(d, "losses.f32"), dtype=np.float32).astype(np.float64))
dev = torch.device("cuda:0")
rng0 = O.Pcg32(seed)
batches = []
for i in range(steps):
    x_np = O.random_uniform(rng0, B * n_in).reshape(B, n_in)
    batches.append((torch.from_numpy(x_np).to(dev), torch.from_numpy(O.synth_target(x_np, n_out)).to(dev)))
for rep in range(4):
    tr = tcnn.create_from_config(n_in, n_out, cfg, seed=1337, device=dev).trainer
    if rep >= 2: tr.use_fused_step = False
    losses = torch.zeros(steps, device=dev)
    for i, (x, t) in enumerate(batches):
        tr.training_step(x, t); losses[i] = tr.loss_value[0]
    show(f"ours{rep} fused={tr.use_fused_step}", losses.cpu().numpy().astype(np.float64))
