"""Multi-GPU parity check, run under torchrun (tests/test_gpu_parity.py::test_peer_memory_step_on_all_gpus):
peer-memory data parallelism (fused reduce-scatter + Adam + all-gather kernel, P2P and NVSwitch-multicast variants) must
produce the parameters of the NCCL all-reduce path, and every rank must end up with identical fp16 parameters."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tiny-cuda-nn_b200")]
from tests import common  # noqa: E402


def main():
    import tinycudann as tcnn
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    B, steps = 1 << 14, 4
    # DP_CHECK_CONFIG=wide_dp: BASELINE config #5's shape (3-D grid + 128 x 3 network, unfused scatter) instead of config #1
    CFG, n_in, n_out = (common.CONFIG_WIDE_DP, 3, 3) if os.environ.get("DP_CHECK_CONFIG") == "wide_dp" else (common.CONFIG_1, 2, 3)
    ref = tcnn.create_from_config(n_in, n_out, CFG, seed=1337, device=dev, dp_mode="allreduce").trainer
    p0 = ref.params_full_precision.clone()
    batches = []
    for i in range(steps):
        x_np, t_np = common.synth_inputs(B * world, n_in, n_out, seed=70 + i)
        s, e = tcnn.shard_bounds(B * world, rank, world)
        batches.append((torch.from_numpy(x_np[s:e]).to(dev), torch.from_numpy(t_np[s:e]).to(dev)))
    for x, t in batches:
        ref.training_step(x, t)
    ref_loss = ref.loss()
    d_ref = (ref.params_full_precision - p0).double()
    for mc in (False, True):
        tr = tcnn.create_from_config(n_in, n_out, CFG, seed=1337, device=dev, dp_mode="peer", use_multicast=mc).trainer
        if mc and not tr.peer_memory.multicast:
            if rank == 0:
                print("no multicast mapping on this box: NVSwitch variant skipped")
            continue
        for x, t in batches:
            tr.training_step(x, t)
        loss = tr.loss()
        assert abs(loss - ref_loss) <= 2e-3 * abs(ref_loss), (loss, ref_loss)
        full = tr.gather_params_full_precision()
        d = (full - p0).double()
        rel = float((d - d_ref).norm() / d_ref.norm())
        assert rel < 3e-2, f"peer-memory update differs from the all-reduce update: {rel}"  # fp16 atomics order + fp32 vs fp16 reduction
        # every rank holds the same fp16 parameters, and they are the cast of the gathered master weights
        mine = tr.params.clone()
        other = mine.clone()
        dist.broadcast(other, src=0)
        assert torch.equal(mine, other), "fp16 parameters differ between ranks"
        assert torch.equal(mine, full.half()), "fp16 parameters are not the cast of the master weights"
        assert not tr.param_gradients[int(tr.info.n_params_mlp):].any().item(), "grid gradients not re-zeroed"
        torch.cuda.synchronize()
        dist.barrier()
        if rank == 0:
            print(f"multicast={mc}: loss {loss:.6e} (all-reduce path {ref_loss:.6e}), update rel diff {rel:.2e}")
    if rank == 0:
        print("dp_peer_check ok")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
