"""Development bench (GPU box): small-batch sweep of the fused training step with and without the static CUDA graph (a17), beside the
reference build's own number (which runs its step under its CudaGraph capture, cuda_graph.h:65-153).

    python tests/dev/graph_sweep.py [cfg1|oneblob] > profiles/r02_graph_sweep.jsonl

Per batch size: device time per step (CUDA events around `steps` back-to-back calls, which is host-bound when the host cannot keep up)
and the host time per call (wall clock of the Python loop without a sync)."""
import json
import os
import subprocess
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tiny-cuda-nn_b200")]
from tests import common  # noqa: E402
import tinycudann as tcnn  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
which = sys.argv[1] if len(sys.argv) > 1 else "cfg1"
cfg = {"cfg1": common.CONFIG_1, "oneblob": common.CONFIG_ONEBLOB, "oneblob64": common.CONFIG_ONEBLOB_64}[which]
dev = torch.device("cuda:0")
n_in, n_out = 2, 3
steps, warmup = 500, 50

model = tcnn.create_from_config(n_in, n_out, cfg, device=dev)
tr = model.trainer
for k in range(10, 19):
    B = 1 << k
    xs = [torch.rand(B, n_in, device=dev) for _ in range(4)]
    ts = [torch.rand(B, n_out, device=dev) for _ in range(4)]
    row = {"config": which, "B": B}
    for mode in (0, 1):
        tr.set_graph_mode(mode)
        for i in range(warmup):
            tr.training_step(xs[i % 4], ts[i % 4])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        for i in range(steps):
            tr.training_step(xs[i % 4], ts[i % 4])
        t1 = time.perf_counter()
        e1.record(); torch.cuda.synchronize()
        tag = "graph" if mode else "plain"
        row[tag + "_ms_per_step"] = round(e0.elapsed_time(e1) / steps, 5)
        row[tag + "_host_us_per_call"] = round((t1 - t0) / steps * 1e6, 2)
    row["graph_stats"] = tr.graph_stats()
    if os.path.exists(REF):
        with tempfile.TemporaryDirectory() as d:
            json.dump(cfg, open(os.path.join(d, "cfg.json"), "w"))
            r = subprocess.run([REF, "bench", os.path.join(d, "cfg.json"), str(n_in), str(n_out), str(B), "200", "20"], capture_output=True, text=True, timeout=600)
            try:
                row["reference_ms_per_step"] = round(json.loads(r.stdout.strip().splitlines()[-1])["train_ms_per_step"], 5)
            except Exception:  # noqa: BLE001
                row["reference_error"] = (r.stderr or r.stdout)[-200:]
    print(json.dumps(row), flush=True)
