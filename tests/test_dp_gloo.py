"""CPU, world_size 2 over gloo: the data-parallel host logic (sharding, gradient all-reduce, 1/world folding) and the
DP semantics checked with the oracle: sum of the per-shard loss-scaled gradients / world == gradient of the global batch."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import tcnn_oracle as O
from tests import common


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _grads(g, m, p16, x, t, n_out):
    enc = np.zeros((x.shape[0], m.n_input), np.float16)
    enc[:, : g.n_features] = O.grid_forward(g, x, p16[m.n_params:])
    y, hidden = O.mlp_forward(m, p16[: m.n_params], enc)
    lv, dy = O.loss_l2(y, t, n_out)
    dW, dx = O.mlp_backward(m, p16[: m.n_params], enc, hidden, dy)
    return lv, np.concatenate([dW, O.grid_backward(g, x, dx[:, : g.n_features])])


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tinycudann import parallel
        cfg, n_in, n_out, B = common.CONFIG_SHIPPED, 2, 3, 1024
        g, m = common.oracle_configs(cfg, n_in, n_out)
        p16 = O.init_params(m, g, 1337, True).astype(np.float16)
        x, t = common.synth_inputs(B, n_in, n_out, seed=4)
        lo, hi = parallel.shard_bounds(B, rank, world)
        assert (hi - lo) % 256 == 0 and hi - lo == B // world
        lv, grad = _grads(g, m, p16, x[lo:hi], t[lo:hi], n_out)
        gt = torch.from_numpy(grad.astype(np.float16))
        for h in parallel.allreduce_gradients(gt):
            h.wait()
        loss = torch.tensor([lv]); dist.all_reduce(loss)
        if rank == 0:
            full_lv, full = _grads(g, m, p16, x, t, n_out)
            got = gt.numpy().astype(np.float64) / world  # the 1/world the Adam kernel applies via grad_divisor
            out.put((common.rel_l2(got[: m.n_params], full[: m.n_params]), common.rel_l2(got[m.n_params:], full[m.n_params:]),
                     abs(loss.item() / world - full_lv) / full_lv, parallel.grad_divisor(128.0, world)))
    finally:
        dist.destroy_process_group()


def test_dp_gradient_allreduce_equals_global_batch():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    e_mlp, e_grid, e_loss, div = q.get(timeout=10)
    assert e_mlp < 5e-3 and e_grid < 5e-3 and e_loss < 1e-5 and div == 256.0


def test_shard_bounds_validation():
    from tinycudann import parallel
    assert parallel.shard_bounds(1 << 18, 3, 8) == (3 << 15, 4 << 15)
    with pytest.raises(ValueError):
        parallel.shard_bounds(1000, 0, 2)
