"""Development aid: stage timestamps of dp_adam_kernel (build with -DDP_STAMPS) on N ranks."""
import os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tiny-cuda-nn_b200")]
from tests import common
import tinycudann as tcnn
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", rank); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
mc = os.environ.get("MC", "0") == "1"
tr = tcnn.create_from_config(2, 3, common.CONFIG_1, seed=1337, device=dev, dp_mode="peer", use_multicast=mc).trainer
tr.peer_memory.comm.reserved = int(os.environ.get("VAR", "0"))
B = 1 << 18
x = torch.rand(B, 2, device=dev); t = torch.rand(B, 3, device=dev)
acc = torch.zeros(3, dtype=torch.float64)
n = 0
for i in range(60):
    tr.training_step(x, t)
    if i >= 20 and i % 4 == 0:
        torch.cuda.synchronize()
        st = tr.peer_memory.flags.view(torch.int64)[12:16].cpu()
        acc += torch.tensor([st[1] - st[0], st[2] - st[1], st[3] - st[2]], dtype=torch.float64); n += 1
torch.cuda.synchronize(); dist.barrier()
if os.environ.get("ONLY0") and rank != 0: sys.exit(0)
print(f"[r{rank}] mc={mc} var={os.environ.get("VAR")} entry-wait {acc[0]/n/1e3:.1f} us, body {acc[1]/n/1e3:.1f} us, exit-wait {acc[2]/n/1e3:.1f} us", flush=True)
dist.destroy_process_group()
