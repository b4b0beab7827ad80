"""Experiment (GPU box): time grid forward / backward per group of levels to locate scatter contention."""
import os, sys, copy
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tiny-cuda-nn_b200"))
import torch
import tinycudann as tcnn
from tinycudann import _lib
from tinycudann._lib import check, ptr
from tests import common

dev = torch.device("cuda:0")
B = 1 << 18
S = lambda: _lib.C.c_void_p(torch.cuda.current_stream().cuda_stream)
x = torch.rand(B, 2, device=dev)
def timeit(fn, n=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for base, nl in [(16, 4), (256, 4), (4096, 4), (65536, 4), (16, 1), (32, 1), (64, 1), (128, 1), (256, 1), (512, 1), (1024, 1), (2048, 1), (16, 16)]:
    cfg = copy.deepcopy(common.CONFIG_1)
    cfg["encoding"]["base_resolution"] = base; cfg["encoding"]["n_levels"] = nl
    nat = tcnn.NativeModule(2, 3, cfg)
    info = nat.info
    E = info.encoded_width
    table = (torch.rand(info.n_params_grid, device=dev) - 0.5).half()
    enc = torch.empty(B, E, dtype=torch.half, device=dev)
    dy = (torch.randn(B, E, device=dev) * 0.01).half()
    grads = torch.zeros(info.n_params_grid, dtype=torch.half, device=dev); ws = nat.workspace(B, dev)
    tf = timeit(lambda: check(nat.lib.tcnn_b200_encode_forward(nat.handle, S(), B, ptr(x), ptr(table), ptr(enc))))
    tb = timeit(lambda: check(nat.lib.tcnn_b200_encode_backward(nat.handle, S(), B, ptr(x), ptr(dy), ptr(grads), 0, ptr(ws))))
    print(f"base={base:6d} levels={nl:2d} E={E} entries={info.n_params_grid//2:9d}  fwd {tf:8.1f} us  bwd {tb:8.1f} us", flush=True)
