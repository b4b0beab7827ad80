"""Data-parallel plumbing (one process per GPU, torch.distributed).  The reference is single-GPU; the
hot path shards by batch: every rank holds a full replica of the parameters and optimizer state, trains on its own
slice of the global batch, and the ranks exchange exactly one thing per step: the sum of the loss-scaled fp16
gradient buffer [MLP | grid].  Loss gradients are normalised by the LOCAL batch (losses/l2.h:62-64), so the global-batch
gradient is sum / world_size; that 1/world_size is folded into the Adam kernel's gradient divisor."""
from __future__ import annotations

import torch
import torch.distributed as dist

BATCH_SIZE_GRANULARITY = 256


def shard_bounds(global_batch: int, rank: int, world: int, granularity: int = BATCH_SIZE_GRANULARITY):
    """Contiguous [start, end) slice of the global batch for `rank`; every shard is a multiple of `granularity`."""
    if global_batch % (world * granularity) != 0:
        raise ValueError(f"global batch {global_batch} must be a multiple of world_size * {granularity} = {world * granularity}")
    per = global_batch // world
    return rank * per, (rank + 1) * per


def grad_divisor(loss_scale: float, world: int) -> float:
    return loss_scale * world


def allreduce_gradients(grads: torch.Tensor, group=None, bucket_elems: int | None = None):
    """In-place SUM all-reduce of the gradient buffer, optionally in buckets (async handles are returned so that the
    caller can overlap the optimizer on finished buckets).  NCCL reduces fp16 natively; gloo (CPU tests) has no fp16
    arithmetic, so the reduction goes through fp32 there."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return []
    if dist.get_backend(group) == "gloo" and grads.dtype == torch.half:
        tmp = grads.float()
        dist.all_reduce(tmp, group=group)
        grads.copy_(tmp.half())
        return []
    if bucket_elems is None or bucket_elems >= grads.numel():
        return [dist.all_reduce(grads, group=group, async_op=True)]
    handles = []
    for start in range(0, grads.numel(), bucket_elems):
        handles.append(dist.all_reduce(grads[start:start + bucket_elems], group=group, async_op=True))
    return handles


class PeerMemory:
    """Symmetric fp16 parameter / gradient buffers + barrier flags for tcnn_b200_training_step_dp: every rank allocates the
    same sizes from torch's symmetric-memory allocator, the rendezvous maps all ranks' buffers (and, on NVSwitch systems,
    a multicast alias) into each process, and the pointers go into a tcnn_b200_dp_comm."""

    def __init__(self, native, device, group=None, use_multicast: bool | None = None):
        import torch.distributed._symmetric_memory as symm

        from . import _lib
        lib = native.lib
        group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > 8:
            raise RuntimeError("peer-memory data parallelism spans one NVSwitch domain (at most 8 ranks)")
        n = int(native.info.n_params)
        self.native = native
        n_flags = int(lib.tcnn_b200_dp_flag_bytes()) // 4
        if self.world == 1:  # nothing to map: plain allocations, the rank is its own only peer
            self.params = torch.zeros(n, dtype=torch.half, device=device)
            self.grads = torch.zeros(n, dtype=torch.half, device=device)
            self.flags = torch.zeros(n_flags, dtype=torch.int32, device=device)
            ptrs = [([t.data_ptr()], 0) for t in (self.params, self.grads, self.flags)]
        else:
            self.params = symm.empty(n, dtype=torch.half, device=device)
            self.grads = symm.empty(n, dtype=torch.half, device=device)
            self.flags = symm.empty(n_flags, dtype=torch.int32, device=device)
            for t in (self.params, self.grads, self.flags):
                t.zero_()
            self.handles = [symm.rendezvous(t, group.group_name) for t in (self.params, self.grads, self.flags)]
            ptrs = [(list(h.buffer_ptrs), int(h.multicast_ptr or 0)) for h in self.handles]
        (pp, pmc), (gp, gmc), (fp, _) = ptrs
        c = _lib.DpComm()
        c.rank, c.world_size, c.epoch = self.rank, self.world, 0
        for r in range(self.world):
            c.peer_params[r], c.peer_grads[r], c.peer_flags[r] = pp[r], gp[r], fp[r]
        if use_multicast is None:
            # measured (profiles/r01d_bench_n*): 2 ranks 0.235 ms/step with peer loads vs 0.247 through the switch;
            # 4 ranks 0.240 vs 0.235; 8 ranks 0.248 vs 0.237
            use_multicast = self.world > 2
        self.multicast = bool(use_multicast and pmc and gmc)
        if self.multicast:
            c.params_multicast, c.grads_multicast = pmc, gmc
        self.comm = c
        self.shard = self.shard_of(self.rank)
        torch.cuda.synchronize(device)
        dist.barrier(group)  # every rank's zero-fill has landed before anyone signals into the flags

    def shard_of(self, rank: int):
        import ctypes as C
        b, e = C.c_uint64(), C.c_uint64()
        from ._lib import check
        check(self.native.lib.tcnn_b200_dp_shard(self.native.handle, rank, self.world, C.byref(b), C.byref(e)))
        return int(b.value), int(e.value)
