"""Development bench (GPU box, under torchrun): BASELINE config #5 - data-parallel training of a 3-D HashGrid (T = 2^22, 16 x 2)
+ FullyFusedMLP 128 x 3, STRONG scaling of a global batch of 2^22 samples over the ranks (peer-memory step: fused
reduce-scatter + sharded Adam + all-gather kernel).  Device time, max over ranks; rank 0 prints one JSON line.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tests/dev/bench_dp_cfg5.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tiny-cuda-nn_b200")]
from tests import common  # noqa: E402
import tinycudann as tcnn  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = json.loads(json.dumps(common.CONFIG_WIDE_DP)); cfg["encoding"]["log2_hashmap_size"] = 22
    G = 1 << int(os.environ.get("LOG2_GLOBAL_BATCH", "22"))
    B = G // world
    steps, warmup = int(os.environ.get("STEPS", "20")), 5
    model = tcnn.create_from_config(3, 3, cfg, device=dev, dp_mode=os.environ.get("DP_MODE", "peer" if world > 1 else "allreduce"))
    xs = [torch.rand(B, 3, device=dev) for _ in range(2)]
    ts = [torch.rand(B, 3, device=dev) for _ in range(2)]
    for i in range(warmup):
        model.trainer.training_step(xs[i % 2], ts[i % 2])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        model.trainer.training_step(xs[i % 2], ts[i % 2])
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    loss = model.trainer.loss()
    if rank == 0:
        print(json.dumps({"case": "cfg5_T22_128x3_global_b%d" % (G.bit_length() - 1), "n_gpus": world, "global_batch": G, "per_gpu_batch": B, "ms_per_step": float(ms.item()),
                          "samples_per_s": G / float(ms.item()) * 1e3, "scaling": "strong", "loss": loss, "n_params": int(model.trainer.n_params)}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
