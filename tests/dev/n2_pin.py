"""Development aid (GPU box): pins the snapshot format against the reference build in both directions.

  python tests/dev/n2_pin.py gpurun_out/n2

1. `ref_driver serialize`: the reference trains N steps, writes json::to_msgpack(Trainer::serialize(true)) and the losses of its next
   two steps.  -> tests/golden/snapshot_small_2d.msgpack + snapshot_small_2d.json (copied by hand from the output directory).
2. our Trainer.deserialize of that file, then the same two steps: losses printed next to the reference's.
3. our Trainer trains the same N steps, serializes; `ref_driver deserialize` reads OUR file and runs the two steps."""
import json
import os
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tiny-cuda-nn_b200")]
import msgpack  # noqa: E402

import tinycudann as tcnn  # noqa: E402
from oracle import tcnn_oracle as O  # noqa: E402
from tests import common  # noqa: E402

out = os.path.abspath(sys.argv[1]); os.makedirs(out, exist_ok=True)
REF = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
cfg = common.CONFIG_SNAPSHOT
n_in, n_out, B, seed, N = 2, 3, 4096, 77, 20
cfg_path = os.path.join(out, "cfg.json"); json.dump(cfg, open(cfg_path, "w"))
snap = os.path.join(out, "snapshot_small_2d.msgpack")
r = subprocess.run([REF, "serialize", cfg_path, str(n_in), str(n_out), str(B), str(seed), str(N), snap, out], capture_output=True, text=True)
print(r.stdout, r.stderr); assert r.returncode == 0
ref_losses = np.fromfile(os.path.join(out, "losses.f32"), dtype=np.float32)
json.dump({"config": cfg, "n_in": n_in, "n_out": n_out, "B": B, "seed": seed, "n_steps": N, "losses": [float(v) for v in ref_losses]},
          open(os.path.join(out, "snapshot_small_2d.json"), "w"), indent=1)

dev = torch.device("cuda:0")


def batches(skip, n):
    rng = O.Pcg32(seed)
    for i in range(skip + n):
        x = O.random_uniform(rng, B * n_in).reshape(B, n_in)
        if i >= skip:
            yield torch.from_numpy(x).to(dev), torch.from_numpy(O.synth_target(x, n_out)).to(dev)


# 2. reference snapshot -> our trainer
tr = tcnn.create_from_config(n_in, n_out, cfg, seed=5, device=dev).trainer
tr.deserialize(msgpack.unpackb(open(snap, "rb").read(), raw=False))
ours = []
for x, t in batches(N, 2):
    tr.training_step(x, t); ours.append(tr.loss())
print("ref snapshot -> ours: next losses", ours, "reference", ref_losses[N:].tolist(), "step", tr.current_step)

# 3. our snapshot -> reference
tr = tcnn.create_from_config(n_in, n_out, cfg, seed=1337, device=dev).trainer
mine = []
for x, t in batches(0, N):
    tr.training_step(x, t); mine.append(tr.loss())
print("our first", N, "losses vs reference: max rel", float(np.max(np.abs(np.array(mine) - ref_losses[:N]) / ref_losses[:N])))
our_snap = os.path.join(out, "ours.msgpack")
open(our_snap, "wb").write(msgpack.packb(tr.serialize(serialize_optimizer=True), use_bin_type=True))
nxt = []
for x, t in batches(N, 2):
    tr.training_step(x, t); nxt.append(tr.loss())
r = subprocess.run([REF, "deserialize", cfg_path, str(n_in), str(n_out), str(B), str(seed), str(N), our_snap], capture_output=True, text=True)
print(r.stdout, r.stderr); assert r.returncode == 0
print("our snapshot -> reference: reference next", json.loads(r.stdout.strip().splitlines()[-1])["next_losses"], "ours", nxt)
