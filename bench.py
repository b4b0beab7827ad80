#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (BASELINE.json): training samples/sec of
HashGrid(2-D, 16 levels x 2 features, T=2^19) + FullyFusedMLP(64 x 2) + L2 + Adam at batch 2^18 per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One process per GPU (torchrun for N > 1); each rank trains on its own 2^18-sample shard (weak scaling) and the fp16
gradient buffer is all-reduced once per step.  Prints ONE JSON line on rank 0.  See DESIGN.md section "Measurement"."""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tiny-cuda-nn_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

BATCH = 1 << 18
N_IN, N_OUT = 2, 3
METRIC = "training samples/sec HashGrid+FullyFusedMLP64 @ batch 2^18, 1/2/4/8 B200"
WORKLOAD = "cfg1: 2D HashGrid 16x2 T=2^19 base16 scale2.0 + FullyFusedMLP 64x2 ReLU, L2, Adam, batch 2^18 per GPU (README.md:19-42 config of the reference)"
CONFIG_NAME = "cfg1"


def select_workload(name):
    """--config: cfg1 is the metric's configuration (the default, the only one the driver runs); the others are BASELINE.json's
    secondary configurations, same JSON line, `metric` saying which."""
    global BATCH, N_IN, N_OUT, METRIC, WORKLOAD, CONFIG_NAME
    from tests import common
    CONFIG_NAME = name
    if name == "cfg1":
        return common.CONFIG_1
    if name == "cfg2":  # BASELINE.json configs[1]: benchmarks/image/bench_ours.cu with data/config_oneblob.json at 128 neurons
        BATCH, N_IN, N_OUT = 1 << 18, 2, 3
        METRIC = "training samples/sec OneBlob+FullyFusedMLP128x5 @ batch 2^18 (bench_ours.cu)"
        WORKLOAD = "cfg2: OneBlob 64 bins (2-D -> 128 features) + FullyFusedMLP 128x5 ReLU, RelativeL2, Adam, batch 2^18 per GPU (data/config_oneblob.json)"
        return common.CONFIG_ONEBLOB
    if name == "cfg3":  # configs[2]
        BATCH, N_IN, N_OUT = 1 << 20, 3, 4
        METRIC = "training samples/sec NeRF-style 3D HashGrid+FullyFusedMLP64 @ batch 2^20"
        WORKLOAD = "cfg3: 3D HashGrid 16x2 T=2^19 scale1.38 + FullyFusedMLP 64x2, L2, Adam, batch 2^20 per GPU"
        return common.CONFIG_NERF
    if name in ("cfg4", "cfg5"):
        base, log2_T, b = (common.CONFIG_WIDE_F4, 24, 20) if name == "cfg4" else (common.CONFIG_WIDE_DP, 22, 19)
        c = json.loads(json.dumps(base)); c["encoding"]["log2_hashmap_size"] = log2_T
        BATCH, N_IN, N_OUT = 1 << b, 3, (4 if name == "cfg4" else 3)
        METRIC = f"training samples/sec 3D HashGrid T=2^{log2_T}+FullyFusedMLP128 @ batch 2^{b}"
        WORKLOAD = (f"{name}: 3D HashGrid 16x{c['encoding']['n_features_per_level']} T=2^{log2_T} + FullyFusedMLP 128x{c['network']['n_hidden_layers']}, "
                    f"L2, Adam, batch 2^{b} per GPU")
        return c
    raise SystemExit(f"unknown --config {name}")


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


def load_l2_peaks():
    """profiles/l2_peaks.json: random-address reduction / load rates of one B200's L2, measured by tests/dev/l2_peaks.cu."""
    try:
        with open(os.path.join(ROOT, "profiles", "l2_peaks.json")) as f:
            d = json.load(f)
        return {"red": float(d["red_gops_per_s"]["table_32MiB_16B"]), "ld": float(d["ld_gops_per_s"]["table_32MiB_16B"])}
    except Exception:  # noqa: BLE001
        return None


def l2_rooflines(info, B, kernel_ms, launches):
    """The L2-bound kernels against the L2's own ceilings.  Operations per launch: a corner pair (the two
    dim-0 neighbours) costs one 16-byte slot access plus a second access when the pair straddles a slot (1 time in 4), i.e.
    1.25 x 2^(D-1) per sample and level - loads in the gather, red.global.add.f16x2 in the scatter.  When the fused kernel gathers
    its own rows (no grid_forward launch) it carries both; loads and reductions share one L2 budget (profiles/l2_peaks.json,
    "mixed"), so its floor is the SUM of the two times and frac = floor / measured."""
    peaks = load_l2_peaks()
    if not peaks or not info.n_levels or info.n_features_per_level != 2:
        return None
    ops = B * int(info.n_levels) * (1 << (int(info.n_input_dims) - 1)) * 1.25
    out = {"source": "profiles/l2_peaks.json (tests/dev/l2_peaks.cu on a B200)", "unit": "Gop/s",
           "peaks": {"red.global.add.noftz.f16x2 (v1/v4)": peaks["red"], "ld.global.nc 4/16 B": peaks["ld"]}}
    gather_fused = launches.get("grid_forward", 0) == 0
    work = {"mlp_fused": (["red", "ld"] if gather_fused else ["red"]), "grid_forward": ([] if gather_fused else ["ld"])}
    for kernel, kinds in work.items():
        ms = kernel_ms.get(kernel, 0.0)
        if ms > 0 and kinds:
            floor_ms = sum(ops / (peaks[k] * 1e9) * 1e3 for k in kinds)
            out[kernel] = {"ops_per_launch": {k: ops for k in kinds}, "achieved": len(kinds) * ops / (ms * 1e-3) / 1e9, "floor_ms": floor_ms, "ms": ms, "frac": floor_ms / ms}
    return out


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled while the timed region runs (B200_PROFILING.md recipe)."""

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:  # noqa: BLE001
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:  # noqa: BLE001
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:  # noqa: BLE001
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def algorithmic_bytes(info, B):
    """A_train = B(4D+4O) + 38 P + sum_levels min(B 2^D 32, 2 F N_level), and per-kernel shares."""
    D, O, F, L = info.n_input_dims, info.n_output_dims, info.n_features_per_level, info.n_levels
    gathers = sum(min(B * (1 << D) * 32, 2 * F * (info.level_offset[l + 1] - info.level_offset[l])) for l in range(L))
    io = B * (4 * D + 4 * O)
    P = int(info.n_params)
    per_kernel = {
        "grid_forward": B * 4 * D + gathers + B * 2 * info.encoded_width,  # coordinates + every touched table sector once + encoded out
        # fused MLP + loss + backward + scatter: encoded batch in (fp16), targets, coordinates, weights, one pass over the
        # gradient table (activations and dL/d(encoded) never leave the SM)
        "mlp_fused": B * (2 * info.encoded_width + 4 * O + 4 * D) + 2 * int(info.n_params_mlp) + 2 * int(info.n_params_grid),
        "mlp_finalize": 2 * int(info.n_params_mlp),
        "grid_backward": 0,  # only the replica memset / reduce of the coarse levels remains a separate launch
        "adam": 38 * P - 2 * int(info.n_params_grid),          # 38 B/param minus the gradient write counted in grid_backward
    }
    per_kernel["mlp_fused_with_gather"] = B * (4 * O + 4 * D) + gathers + 2 * int(info.n_params_mlp) + 2 * int(info.n_params_grid)
    return io + 38 * P + gathers, per_kernel


def cpu_baseline_port(cfg, seconds_budget=12.0):
    """Pure-PyTorch CPU port (oracle/torch_cpu_port.py) on all host cores, bounded sample of the same workload."""
    import numpy as np
    import torch

    from oracle import tcnn_oracle as O
    from oracle import torch_cpu_port as port
    from tests import common

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    g, m = common.oracle_configs(cfg, N_IN, N_OUT)
    params = O.init_params(m, g, 1337, True)
    o = cfg["optimizer"]
    _, step = port.make_trainer(g, m, params, O.AdamConfig(o.get("learning_rate", 1e-3)))
    Bc = 1 << 15
    x_np, t_np = common.synth_inputs(Bc, N_IN, N_OUT, seed=5)
    x, t = torch.from_numpy(x_np), torch.from_numpy(t_np)
    step(x, t)  # warm-up
    n, t0 = 0, time.perf_counter()
    while True:
        step(x, t); n += 1
        el = time.perf_counter() - t0
        if el > seconds_budget or n >= 50:
            break
    return {"value": Bc * n / el, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": f"{n} training steps of batch 2^15 (1/8 of the GPU batch), torch {torch.__version__} fp32, {cores} threads"}


def run_reference(args, cfg):
    """Reference arm.  The reference has no CPU path, so — as BASELINE.json prescribes — the arm is the
    reference's own CUDA build (oracle/_ref/ref_driver: unmodified reference sources + our small harness over its
    public C++ API) on one B200, same config and batch; the pure-PyTorch CPU port is reported as cpu_baseline."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ref = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    line = {"impl": "reference", "metric": METRIC, "unit": "samples/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": config_dict(args.gpus), "reference_note": "the reference is single-GPU: its 1-GPU number is reported for every --gpus"}
    try:
        cpu = cpu_baseline_port(cfg)
    except Exception as e:  # noqa: BLE001
        cpu = {"unavailable": str(e)[:200], "value": None}
    have_gpu = False
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:  # noqa: BLE001
        pass
    if os.path.exists(ref) and have_gpu:
        import tempfile
        with tempfile.TemporaryDirectory() as d:
            cp = os.path.join(d, "cfg.json")
            json.dump(cfg, open(cp, "w"))
            steps = min(args.steps, 1000)
            r = subprocess.run([ref, "bench", cp, str(N_IN), str(N_OUT), str(BATCH), str(steps), str(max(args.warmup, 3))], capture_output=True, text=True, timeout=900)
        if r.returncode == 0:
            j = json.loads(r.stdout.strip().splitlines()[-1])
            line.update({"value": j["train_samples_per_s"], "ms_per_step": j["train_ms_per_step"], "steps": steps,
                         "e2e": {"value": j["train_e2e_samples_per_s"], "unit": "samples/s", "h2d_bytes_per_step": BATCH * 4 * (N_IN + N_OUT), "d2h_bytes_per_step": 4,
                                 "ms_per_step": j["train_e2e_ms_per_step"],
                                 "how": "pinned host batch -> H2D on a second stream into the other of two device buffers under the previous step -> Trainer::training_step -> Trainer::loss (the reference's blocking loss read-back) every step"},
                         "reference_kind": "reference CUDA build (nvcc -gencode arch=compute_100,code=sm_100 of /root/reference/src), 1 B200",
                         "inference_samples_per_s": j["inference_samples_per_s"], "cpu_baseline": cpu, "gpu_launches": None})
            print(json.dumps(line), flush=True)
            return
        line["reference_error"] = (r.stderr or r.stdout)[-300:]
    # no GPU / no binary: the CPU port is the only thing that can run
    line.update({"value": cpu["value"], "ms_per_step": BATCH / cpu["value"] * 1e3, "cpu_baseline": cpu,
                 "e2e": {"value": cpu["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "reference_kind": "pure-PyTorch CPU port (reference CUDA build unavailable here)"})
    print(json.dumps(line), flush=True)


def bind_to_gpu_numa_node(local_rank):
    """Run this rank's host threads on the CPUs next to its GPU before any pinned buffer is allocated (first touch places the pages on
    that NUMA node): at N >= 4 every rank pulls 5.2 MB per step over PCIe and a buffer on the far socket halves that rate.
    -> description of what was done, for the JSON line."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local_rank)
        if all(hasattr(pr, k) for k in ("pci_domain_id", "pci_bus_id", "pci_device_id")):
            bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        else:
            import pynvml
            pynvml.nvmlInit()
            bdf = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local_rank)).busId
            bdf = (bdf.decode() if isinstance(bdf, bytes) else bdf).lower()
            if len(bdf.split(":")[0]) == 8:  # nvml prints an 8-digit PCI domain, sysfs uses 4
                bdf = bdf[4:]
        path = f"/sys/bus/pci/devices/{bdf}/local_cpulist"
        cpus = set()
        for part in open(path).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return "no local CPU of the GPU in this process's affinity mask"
        os.sched_setaffinity(0, cpus)
        node = open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip()
        return f"host threads bound to the {len(cpus)} CPUs local to GPU {bdf} (NUMA node {node}) before the pinned allocations"
    except Exception as e:  # noqa: BLE001 - containers without sysfs PCI topology: run unbound
        return f"unbound ({type(e).__name__}: {str(e)[:80]})"


def config_dict(world, dp_mode="peer"):
    """The `config` object of the JSON line: identical in both arms (ours and --impl reference) for the same --gpus."""
    l2 = ("no flush: one step streams ~0.45 GB of parameter/optimizer state (> 126 MB L2) and cycles 4 distinct resident batches" if CONFIG_NAME == "cfg1" else
          "no flush: 4 distinct resident batches are cycled; parameters + optimizer state + the per-batch activations of one step exceed the 126 MB L2")
    return {"workload": WORKLOAD, "batch_per_gpu": BATCH, "global_batch": BATCH * world, "parallelism": f"dp{world}", "l2": l2}


DP_NOTE = {
    "peer": "one fused kernel per step: sum of this rank's gradient shard over all ranks (NVSwitch multimem.ld_reduce; plain peer loads at 2 GPUs) -> Adam on the shard -> new fp16 parameters to all ranks (multimem.st / peer stores), both barriers in-kernel",
    "peer-mc": "one fused kernel per step: NVSwitch multimem.ld_reduce of this rank's gradient shard -> Adam on the shard -> multimem.st of the new fp16 parameters to all ranks, both barriers in-kernel",
    "peer-p2p": "one fused kernel per step: peer loads of this rank's gradient shard from all ranks -> Adam on the shard -> peer stores of the new fp16 parameters, both barriers in-kernel",
    "allreduce": "fp16 NCCL all-reduce of the 22.4 MB gradient buffer per step, replicated Adam",
}


def run_dp_check(cfg, dev, rank, world, dp_mode):
    """Correctness of the multi-rank step, printed with the bench line: (1) after 50 data-parallel steps on a fixed global batch
    sequence every rank holds bit-identical fp16 parameters; (2) the loss of those steps equals what ONE rank gets on the same
    global batches (run live on rank 0 with a one-rank trainer), up to the summation order of fp16 gradients."""
    import hashlib

    import numpy as np
    import torch
    import torch.distributed as dist

    import tinycudann as tcnn
    from tests import common
    from tinycudann import parallel

    steps, gb = 50, 1 << 16
    solo_group = dist.new_group([0])  # collective: every rank creates it, rank 0 trains alone in it
    kw = {"dp_mode": dp_mode.split("-")[0], "use_multicast": {"peer": None, "peer-p2p": False, "peer-mc": True}.get(dp_mode)}
    tr = tcnn.create_from_config(N_IN, N_OUT, cfg, seed=1337, device=dev, **kw).trainer
    lo, hi = parallel.shard_bounds(gb, rank, world)
    losses = []
    for s in range(steps):
        x_np, t_np = common.synth_inputs(gb, N_IN, N_OUT, seed=500 + s)
        tr.training_step(torch.from_numpy(x_np[lo:hi]).to(dev), torch.from_numpy(t_np[lo:hi]).to(dev))
        if s >= steps - 5:
            losses.append(tr.loss())  # mean over ranks = loss of the global batch
    torch.cuda.synchronize()
    digest = hashlib.sha256(tr.params.cpu().numpy().tobytes()).digest()[:8]
    mine = torch.tensor(list(digest), dtype=torch.int32, device=dev)
    every = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(every, mine)
    identical = all(torch.equal(e, every[0]) for e in every)
    out = None
    if rank == 0:
        one = tcnn.create_from_config(N_IN, N_OUT, cfg, seed=1337, device=dev, process_group=solo_group, dp_mode="allreduce").trainer
        ref = []
        for s in range(steps):
            x_np, t_np = common.synth_inputs(gb, N_IN, N_OUT, seed=500 + s)
            one.training_step(torch.from_numpy(x_np).to(dev), torch.from_numpy(t_np).to(dev))
            if s >= steps - 5:
                ref.append(float(one.loss_value.item()))
        a, b = float(np.mean(losses)), float(np.mean(ref))
        out = {"params_identical": bool(identical), "params_sha256_64": digest.hex(), "steps": steps, "global_batch": gb,
               "loss": a, "loss_1rank": b, "rel": abs(a - b) / b,
               "what": "mean training loss of steps 46-50 on a fixed global batch sequence: N ranks (sharded) vs one rank (whole batch), same seed"}
    dist.barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg1", choices=["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"], help="cfg1 = the metric's configuration (default); cfg2..5 = BASELINE.json's secondary configurations")
    ap.add_argument("--e2e-steps", type=int, default=300)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--force-dp", action="store_true", help="N = 1: run the data-parallel step on a one-rank group")
    ap.add_argument("--dp-mode", default="peer", choices=["peer", "peer-p2p", "peer-mc", "allreduce"],
                    help="N > 1: fused reduce-scatter+Adam+all-gather kernel over NVSwitch multicast (peer) or plain peer loads/stores (peer-p2p), or NCCL all-reduce + replicated Adam (allreduce)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    from tests import common
    cfg = select_workload(args.config)
    if args.impl == "reference":
        return run_reference(args, cfg)

    import torch
    import torch.distributed as dist

    import tinycudann as tcnn

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    numa = bind_to_gpu_numa_node(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    elif args.force_dp:  # development aid: the data-parallel step on a one-rank group (ncu can profile it)
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29533", rank=0, world_size=1, device_id=dev)

    model = tcnn.create_from_config(N_IN, N_OUT, cfg, seed=1337, device=dev, **({} if world == 1 and not args.force_dp else {"dp_mode": args.dp_mode.split("-")[0], "use_multicast": {"peer": None, "peer-p2p": False, "peer-mc": True}.get(args.dp_mode)}))
    tr = model.trainer
    info = tr.info
    lib = tr.lib

    # synthetic data: 4 distinct resident batches per rank, analytic target (tests/common.synth_inputs)
    n_batches = 4
    batches = []
    for b in range(n_batches):
        x_np, t_np = common.synth_inputs(BATCH, N_IN, N_OUT, seed=1000 + 17 * rank + b)
        batches.append((torch.from_numpy(x_np).to(dev), torch.from_numpy(t_np).to(dev)))
    hx = torch.from_numpy(common.synth_inputs(BATCH, N_IN, N_OUT, seed=77 + rank)[0]).pin_memory()
    ht = torch.from_numpy(common.synth_inputs(BATCH, N_IN, N_OUT, seed=77 + rank)[1]).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(n, body):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for i in range(n):
            body(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def train(i):
        x, t = batches[i % n_batches]
        tr.training_step(x, t)

    # End-to-end: every step's inputs come from pinned host memory (H2D inside the timed region) and its loss goes back to the host,
    # through the public call for host-resident batches: Trainer.training_step_host = tcnn_b200_host_batch_begin -> training_step ->
    # tcnn_b200_host_batch_end.  The library copies the NEXT batch on its copy stream under this step's compute and reads the
    # 4-byte loss back on a third stream; the host synchronises once at the end of the timed region.
    hlosses = torch.zeros(4096, dtype=torch.float32).pin_memory()
    E2E_WARMUP = 6  # untimed iterations of the same loop in front: the first, un-overlapped H2D copy is not part of a 20-step measurement

    def train_e2e(i, last=False):
        tr.training_step_host(hx, ht, next_batch=None if last else (hx, ht), host_loss=hlosses[i % 4096: i % 4096 + 1], join=last)

    for i in range(max(args.warmup, n_batches)):  # every resident batch once: its step graph is captured outside the timed region
        train(i)
    with ClockSampler(local_rank) as clk:
        total_ms = timed(args.steps, train)
    ms_per_step = total_ms / args.steps
    value = BATCH * world / (ms_per_step * 1e-3)
    final_loss = tr.loss()

    # Network::inference (object.h:147-176) on the same resident batches: fp32 [B][O] output, device time
    inf_out = torch.empty(BATCH, N_OUT, dtype=torch.float32, device=dev)
    for i in range(5):
        tr.inference(batches[i % n_batches][0], inf_out)
    inf_steps = min(args.steps, 500)
    inf_ms = timed(inf_steps, lambda i: tr.inference(batches[i % n_batches][0], inf_out)) / inf_steps

    e2e_steps = min(args.e2e_steps, args.steps)
    for i in range(E2E_WARMUP):
        train_e2e(i)
    e2e_ms = timed(e2e_steps, lambda i: train_e2e(E2E_WARMUP + i, last=(i == e2e_steps - 1))) / e2e_steps
    h2d_gbs = BATCH * 4 * (N_IN + N_OUT) / (e2e_ms * 1e-3) / 1e9  # what the copy stream must sustain per rank to keep up with the step
    e2e_value = BATCH * world / (e2e_ms * 1e-3)

    # per-kernel device time over a further stretch of the same loop (CUDA events on the launch stream)
    hbm_peak, peak_src = load_peaks()
    a_train, per_kernel_bytes = algorithmic_bytes(info, BATCH)
    kinds = ["grid_forward", "mlp_fused", "mlp_finalize", "grid_backward", "adam"]  # mlp_fused includes the fused gradient scatter
    ms_sum = (ctypes.c_float * 5)(); cnt = (ctypes.c_uint32 * 5)()
    lib.tcnn_b200_profile_enable(tr.native.handle, 1)
    prof_steps = min(200, args.steps)
    for i in range(prof_steps):
        train(i)
    lib.tcnn_b200_profile_read(tr.native.handle, ms_sum, cnt, 5)
    lib.tcnn_b200_profile_enable(tr.native.handle, 0)
    kernel_ms = {k: ms_sum[i] / prof_steps for i, k in enumerate(kinds)}  # device ms per step (all launches of that kind)
    launches_per_step = {k: cnt[i] / prof_steps for i, k in enumerate(kinds)}
    dominant = max(kernel_ms, key=kernel_ms.get)
    dom_ms = kernel_ms[dominant]
    if dominant == "mlp_fused" and launches_per_step.get("grid_forward", 0) == 0 and info.n_levels:
        per_kernel_bytes["mlp_fused"] = per_kernel_bytes["mlp_fused_with_gather"]  # the kernel gathers its own rows: no encoded batch in memory
    achieved = per_kernel_bytes[dominant] / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0  # bytes per step / device time per step of that kernel
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "dram_traffic.json")) as f:
            traffic = json.load(f).get(dominant)
    except Exception:  # noqa: BLE001
        pass

    dp_check = None
    if world > 1:
        dp_check = run_dp_check(cfg, dev, rank, world, args.dp_mode)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16",
            "data": "synthetic",
            "config": config_dict(world), "launch": "one static CUDA graph per (batch, buffers): %d kernel nodes, Adam scalars patched per step" % int(sum(launches_per_step.values())) if world == 1 and tr.graph_stats()[1] else "stream launches", "gradient_exchange": DP_NOTE.get(args.dp_mode, args.dp_mode) if world > 1 else "none",
            "clocks": clk.summary(),
            "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": BATCH * 4 * (N_IN + N_OUT), "d2h_bytes_per_step": 4,
                    "ms_per_step": e2e_ms, "steps": e2e_steps, "warmup": E2E_WARMUP, "h2d_gb_per_s_per_rank": h2d_gbs, "host_placement": numa, "last_loss_on_host": float(hlosses[(E2E_WARMUP + e2e_steps - 1) % 4096]),
                    "how": "Trainer.training_step_host: pinned host batch -> H2D on the library's copy stream (two staging slots, next batch copied under this step) -> training_step (static graph) -> 4-byte loss D2H on a third stream into pinned memory; one host sync at the end"},
            "gpu_launches": int(sum(launches_per_step.values()) * args.steps),
            "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                         "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": per_kernel_bytes[dominant],
                         "kernel_ms_per_step_serial": kernel_ms, "launches_per_step": launches_per_step, "kernel_share": {k: v / max(sum(kernel_ms.values()), 1e-9) for k, v in kernel_ms.items()},
                         "l2": l2_rooflines(info, BATCH, kernel_ms, launches_per_step),
                         "step": {"algorithmic_bytes": a_train, "achieved": a_train / (ms_per_step * 1e-3) / 1e9, "frac": a_train / (ms_per_step * 1e-3) / 1e9 / hbm_peak}},
            "final_loss": final_loss,
            "inference": {"value": BATCH * world / (inf_ms * 1e-3), "unit": "samples/s", "ms_per_batch": inf_ms,
                          "what": f"Network::inference, fp32 output, inputs resident, batch {BATCH} per GPU"},
        }
        if dp_check is not None:
            line["dp_check"] = dp_check
        if world == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_baseline_port(cfg)
            except Exception as e:  # noqa: BLE001 - the CPU port covers grid encodings (the metric's configuration)
                line["cpu_baseline"] = {"unavailable": str(e)[:200]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
