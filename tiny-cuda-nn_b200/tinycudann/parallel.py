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
