"""Development bench (GPU box): training / inference ms per step of OUR library and of the reference build (oracle/_ref/ref_driver
bench) on the secondary BASELINE configs.   python tests/dev/bench_cfg.py [name ...]   -> one JSON line per (config, impl)."""
import json
import os
import subprocess
import sys
import tempfile

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tiny-cuda-nn_b200"))
from tests import common  # noqa: E402
import tinycudann as tcnn  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "ref_driver")


def big(cfg, log2_T):
    c = json.loads(json.dumps(cfg))
    c["encoding"]["log2_hashmap_size"] = log2_T
    return c


CASES = {  # name: (config, n_in, n_out, B)
    "oneblob_128x5_b18": (common.CONFIG_ONEBLOB, 2, 3, 1 << 18),
    "oneblob16_64x4_b18": (common.CONFIG_ONEBLOB16, 2, 3, 1 << 18),
    "oneblob_64x5_b18": (common.CONFIG_ONEBLOB_64, 2, 3, 1 << 18),
    "oneblob_64x5_b20": (common.CONFIG_ONEBLOB_64, 2, 3, 1 << 20),
    "oneblob_128x5_b20": (common.CONFIG_ONEBLOB, 2, 3, 1 << 20),
    "cfg4_T24_128x4_b20": (big(common.CONFIG_WIDE_F4, 24), 3, 4, 1 << 20),
    "cfg5_T22_128x3_b19": (big(common.CONFIG_WIDE_DP, 22), 3, 3, 1 << 19),
    "cfg3_nerf_b20": (common.CONFIG_NERF, 3, 4, 1 << 20),
    # cfg #1 with the reference's narrow network widths
    "cfg1_32x2_b18": (dict(common.CONFIG_1, network=dict(common.CONFIG_1["network"], n_neurons=32)), 2, 3, 1 << 18),
    "cfg1_16x2_b18": (dict(common.CONFIG_1, network=dict(common.CONFIG_1["network"], n_neurons=16)), 2, 3, 1 << 18),
}


def ours(name, cfg, n_in, n_out, B, steps=30, warmup=5):
    dev = torch.device("cuda:0")
    model = tcnn.create_from_config(n_in, n_out, cfg, device=dev)
    xs, ts = [], []
    for i in range(4):
        x = torch.rand(B, n_in, device=dev)
        xs.append(x); ts.append(torch.rand(B, n_out, device=dev))
    out = torch.empty(B, n_out, device=dev)

    def timed(fn):
        for i in range(warmup):
            fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    tr = timed(lambda i: model.trainer.training_step(xs[i % 4], ts[i % 4]))
    inf = timed(lambda i: model.trainer.inference(xs[i % 4], out))
    print(json.dumps({"case": name, "impl": "tcnn_b200", "B": B, "train_ms_per_step": tr, "train_samples_per_s": B / tr * 1e3,
                      "inference_ms_per_step": inf, "inference_samples_per_s": B / inf * 1e3, "last_loss": model.trainer.loss(),
                      "n_params": int(model.trainer.n_params)}), flush=True)
    del model
    torch.cuda.empty_cache()


def reference(name, cfg, n_in, n_out, B, steps=30, warmup=5):
    if not os.path.exists(REF):
        print(json.dumps({"case": name, "impl": "reference", "unavailable": "oracle/_ref/ref_driver not built"}), flush=True)
        return
    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "cfg.json"), "w") as f:
            json.dump(cfg, f)
        r = subprocess.run([REF, "bench", os.path.join(d, "cfg.json"), str(n_in), str(n_out), str(B), str(steps), str(warmup)], capture_output=True, text=True, timeout=600)
        line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else ""
        try:
            j = json.loads(line); j["case"] = name
            print(json.dumps(j), flush=True)
        except Exception:  # noqa: BLE001
            print(json.dumps({"case": name, "impl": "reference", "error": (r.stderr or r.stdout)[-300:]}), flush=True)


# the reference's own benchmark sweep (benchmarks/image/bench_ours.cu:188-331 with data/config_oneblob.json): batch 2^14 .. 2^21
for _k in range(14, 22):
    CASES[f"sweep_oneblob_128x5_b{_k}"] = (common.CONFIG_ONEBLOB, 2, 3, 1 << _k)

if __name__ == "__main__":
    if "--sweep" in sys.argv:
        sys.argv += [n for n in CASES if n.startswith("sweep_")]
    no_ref = "--no-ref" in sys.argv
    names = [a for a in sys.argv[1:] if not a.startswith("--")] or [n for n in CASES if not n.startswith("sweep_")]
    for n in names:
        try:
            ours(n, *CASES[n])
        except Exception as e:  # noqa: BLE001
            print(json.dumps({"case": n, "impl": "tcnn_b200", "error": str(e)[-300:]}), flush=True)
        if not no_ref:
            reference(n, *CASES[n])
