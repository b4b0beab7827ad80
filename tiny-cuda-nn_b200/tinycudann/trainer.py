"""Python mirror of the reference's C++ training facade for the hot path:
`create_from_config` (include/tiny-cuda-nn/config.h:46-63) and `Trainer` (include/tiny-cuda-nn/trainer.h:47-361).

The trainer owns the parameter buffer in the reference's layout (one allocation: fp32 master | fp16 params |
fp16 gradients, trainer.h:75,322-336; MLP weights first, grid tables second), the Adam state
(optimizers/adam.h:128-148), and drives the C-ABI kernels.  Data-parallel training (one process per GPU) adds exactly
one exchange per step: by default a fused reduce-scatter + Adam + all-gather kernel over NVLink peer memory
(dp_mode="peer", tcnn_b200_training_step_dp), else an NCCL all-reduce of the fp16 gradient buffer (dp_mode="allreduce"),
with 1/world_size folded into the Adam kernel either way.
"""
from __future__ import annotations

import json
from dataclasses import dataclass

import torch

from . import _lib
from ._lib import check, ptr
from .modules import NativeModule, _stream_ptr
from .parallel import PeerMemory, allreduce_gradients, grad_divisor

import os

# default_loss_scale<__half>(), common.h:229-233; trainer.h:174.  TCNN_B200_LOSS_SCALE: development override (tests/dev/parity_probe.py)
LOSS_SCALE = float(os.environ.get("TCNN_B200_LOSS_SCALE", "128"))


class Trainer:
    def __init__(self, native: NativeModule, seed: int = 1337, device="cuda", process_group=None, dp_mode: str | None = None, use_multicast: bool | None = None):
        self.native = native
        self.lib = native.lib
        self.device = torch.device(device)
        self.info = native.info
        P = int(self.info.n_params)
        self.n_params = P
        # trainer.h:75: [fp32 master (4P) | fp16 params (2P) | fp16 grads (2P)]
        self.process_group = process_group
        self.world_size = 1
        if process_group is not None or (torch.distributed.is_available() and torch.distributed.is_initialized()):
            self.world_size = torch.distributed.get_world_size(process_group)
        self.peer_memory = None
        if dp_mode is None:
            dp_mode = "peer" if (self.world_size > 1 and self.device.type == "cuda" and torch.distributed.get_backend(process_group) == "nccl") else "allreduce"
        if dp_mode == "peer":
            # fp16 parameters and gradients live in symmetric memory that every rank maps (parallel.PeerMemory)
            # fp32 master weights and Adam state exist for this rank's shard [b, e) only (ZeRO-1 in memory as well as in traffic)
            self.peer_memory = PeerMemory(native, self.device, process_group, use_multicast=use_multicast)
            b, e = self.peer_memory.shard
            self.params_full_precision = torch.zeros(e - b, dtype=torch.float32, device=self.device)
            self.params, self.param_gradients = self.peer_memory.params, self.peer_memory.grads
        else:
            self.params_buffer = torch.zeros(8 * P, dtype=torch.uint8, device=self.device)
            self.params_full_precision = self.params_buffer[: 4 * P].view(torch.float32)
            self.params = self.params_buffer[4 * P : 6 * P].view(torch.half)
            self.param_gradients = self.params_buffer[6 * P : 8 * P].view(torch.half)
        # AdamOptimizer::allocate (adam.h:128-148); peer mode: the shard's state only
        n_state = self.params_full_precision.numel()
        self.first_moments = torch.zeros(n_state, dtype=torch.float32, device=self.device)
        self.second_moments = torch.zeros(n_state, dtype=torch.float32, device=self.device)
        self.param_steps = torch.zeros(n_state, dtype=torch.int32, device=self.device)
        self.current_step = 0
        self.loss_value = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._grid_grads_clean = True  # grid gradient region is all zero (after init / after an optimizer step)
        self.use_fused_step = True     # single GPU: tcnn_b200_training_step (Adam overlapped with the scatter)
        self.initialize_params(seed)

    # trainer.h:68-87
    def initialize_params(self, seed: int = 1337):
        with torch.cuda.device(self.device):
            full = self.params_full_precision if self.peer_memory is None else torch.empty(self.n_params, dtype=torch.float32, device=self.device)
            check(self.lib.tcnn_b200_init_params(self.native.handle, _stream_ptr(), seed, 1, ptr(full)))
            check(self.lib.tcnn_b200_cast_params(_stream_ptr(), self.n_params, ptr(full), ptr(self.params)))
            if self.peer_memory is not None:
                b, e = self.peer_memory.shard
                self.params_full_precision.copy_(full[b:e])
                del full
            self.param_gradients.zero_()
            self.first_moments.zero_(); self.second_moments.zero_(); self.param_steps.zero_()
            self.current_step = 0
            self._grid_grads_clean = True
            torch.cuda.synchronize(self.device)

    def set_params_full_precision(self, params_fp32: torch.Tensor):
        full = params_fp32.to(self.device, torch.float32).contiguous()
        if self.peer_memory is None:
            self.params_full_precision.copy_(full)
        else:
            b, e = self.peer_memory.shard
            self.params_full_precision.copy_(full[b:e])
        check(self.lib.tcnn_b200_cast_params(_stream_ptr(), self.n_params, ptr(full), ptr(self.params)))

    # trainer.h:163-190.  input [B, D] fp32, target [B, O] fp32 (== GPUMatrix CM [D x B] / [O x B]).
    def training_step(self, input: torch.Tensor, target: torch.Tensor, run_optimizer: bool = True, output: torch.Tensor | None = None):
        B = input.shape[0]
        ws = self.native.workspace(B, self.device)
        if self.peer_memory is not None and run_optimizer:
            if not self._grid_grads_clean:
                self.param_gradients[int(self.info.n_params_mlp):].zero_()
                self._grid_grads_clean = True
            return self._peer_step(input, target, output, ws)
        if run_optimizer and self.use_fused_step and self.world_size == 1 and self._grid_grads_clean:
            # single GPU: one fused call, Adam of finished level groups overlaps the remaining gradient scatter
            self.current_step += 1
            check(self.lib.tcnn_b200_training_step(
                self.native.handle, _stream_ptr(), B, ptr(input), ptr(target), LOSS_SCALE, self.current_step,
                ptr(self.params_full_precision), ptr(self.params), ptr(self.param_gradients), ptr(self.first_moments),
                ptr(self.second_moments), ptr(self.param_steps), ptr(self.loss_value), ptr(output), ptr(ws)))
            return self
        check(self.lib.tcnn_b200_training_fwd_bwd(
            self.native.handle, _stream_ptr(), B, ptr(input), ptr(target), ptr(self.params), ptr(self.param_gradients),
            LOSS_SCALE, 0 if self._grid_grads_clean else 1, ptr(self.loss_value), ptr(output), ptr(ws)))
        self._grid_grads_clean = False
        if self.world_size > 1 and self.peer_memory is None:
            for h in allreduce_gradients(self.param_gradients, self.process_group):
                h.wait()  # stream-ordered for NCCL: makes the current stream wait, the host does not block
        if run_optimizer:
            self.optimizer_step(LOSS_SCALE)
        return self

    def _peer_step(self, input, target, output, ws):
        """Data-parallel step over NVLink peer memory: local forward/backward, then ONE kernel that sums this rank's shard
        of the gradients over all ranks, applies Adam to it and stores the new fp16 parameters into every rank
        (tcnn_b200_training_step_dp).  Master weights and Adam state are live only inside the rank's own shard."""
        B = input.shape[0]
        self.current_step += 1
        pm = self.peer_memory
        pm.comm.epoch += 1
        b, e = pm.shard
        check(self.lib.tcnn_b200_training_step_dp(
            self.native.handle, _stream_ptr(), B, ptr(input), ptr(target), LOSS_SCALE, self.current_step,
            ptr(self.params_full_precision), ptr(self.first_moments), ptr(self.second_moments),
            ptr(self.param_steps), ptr(self.loss_value), ptr(output), ptr(ws), _lib.C.byref(pm.comm)))
        return self

    def gather_params_full_precision(self) -> torch.Tensor:
        """fp32 master weights of all shards (collective; in peer-memory mode each rank only updates its own shard)."""
        if self.peer_memory is None:
            return self.params_full_precision.clone()
        full = torch.zeros(self.n_params, dtype=torch.float32, device=self.device)
        b, e = self.peer_memory.shard
        full[b:e] = self.params_full_precision
        for r in range(self.world_size):
            b, e = self.peer_memory.shard_of(r)
            if e > b:
                torch.distributed.broadcast(full[b:e], src=torch.distributed.get_global_rank(self.process_group, r) if self.process_group else r, group=self.process_group)
        return full

    # trainer.h:155-157 + adam.h:150-188
    def optimizer_step(self, loss_scale: float = LOSS_SCALE):
        if self.peer_memory is not None:
            raise RuntimeError("dp_mode='peer' shards the optimizer state: use training_step(run_optimizer=True), or dp_mode='allreduce'")
        self.current_step += 1
        check(self.lib.tcnn_b200_adam_step(
            self.native.handle, _stream_ptr(), grad_divisor(loss_scale, self.world_size), self.current_step,
            ptr(self.params_full_precision), ptr(self.params), ptr(self.param_gradients),
            ptr(self.first_moments), ptr(self.second_moments), ptr(self.param_steps), 1))
        self._grid_grads_clean = True

    def check_peers(self):
        """Data-parallel peer mode: raise if a rank gave up waiting for its peers inside the exchange kernel (the kernel sets an
        error word instead of trapping; TCNN_B200_DP_TIMEOUT_S, default 600 s).  Synchronises."""
        if self.peer_memory is not None:
            word = int(self.lib.tcnn_b200_dp_flag_error_word())
            if int(self.peer_memory.flags[word].item()) != 0:
                raise RuntimeError("tcnn_b200: a data-parallel peer did not reach the gradient exchange in time (TCNN_B200_DP_TIMEOUT_S); "
                                   "this rank's parameters are stale - restart from a snapshot")

    # trainer.h:205-207 (synchronises, like the reference's reduce_sum + D2H)
    def loss(self) -> float:
        self.check_peers()
        if self.world_size > 1:  # mean over the global batch; collective: call on every rank
            v = self.loss_value.clone()
            torch.distributed.all_reduce(v, group=self.process_group)
            return float(v.item()) / self.world_size
        return float(self.loss_value.item())

    # trainer.h:275-315 + adam.h:278-299: the reference's snapshot object (see checkpoint.py)
    def serialize(self, serialize_optimizer: bool = False) -> dict:
        from . import checkpoint
        opt = None
        if serialize_optimizer:
            if self.peer_memory is not None:
                raise RuntimeError("dp_mode='peer' shards the optimizer state over the ranks: serialize the parameters only, or use dp_mode='allreduce'")
            opt = {"current_step": self.current_step, "base_learning_rate": float(self.info.learning_rate),
                   "first_moments": self.first_moments.cpu().numpy(), "second_moments": self.second_moments.cpu().numpy(),
                   "param_steps": self.param_steps.cpu().numpy().view("uint32")}
        return checkpoint.pack(self.params.cpu().numpy(), opt)

    def deserialize(self, data: dict):
        from . import checkpoint
        params, opt = checkpoint.unpack(data)
        if params.size != self.n_params:
            raise RuntimeError(f"Can't set params because CPU buffer has the wrong size: {params.size} != {self.n_params}")
        # float snapshots set the master weights, __half snapshots set both copies from the fp16 values (trainer.h:290-305)
        self.set_params_full_precision(torch.from_numpy(params.astype("float32")))
        if opt is not None:
            if self.peer_memory is not None:
                raise RuntimeError("dp_mode='peer' shards the optimizer state over the ranks")
            self.first_moments.copy_(torch.from_numpy(opt["first_moments"]))
            self.second_moments.copy_(torch.from_numpy(opt["second_moments"]))
            self.param_steps.copy_(torch.from_numpy(opt["param_steps"].view("int32")))
            self.current_step = opt["current_step"]
            self.update_hyperparams({"learning_rate": opt["base_learning_rate"]})
        self.param_gradients.zero_()
        self._grid_grads_clean = True
        torch.cuda.synchronize(self.device)

    # Batches in pinned host memory: H2D of the next batch under this step, loss back on its own stream (tcnn_b200_host_batch_begin/_end).
    def training_step_host(self, host_input: torch.Tensor, host_target: torch.Tensor, next_batch=None, host_loss: torch.Tensor | None = None,
                           join: bool = False, **step_kwargs):
        """One training step on a batch held in (pinned) host memory.  `next_batch` = (host_input, host_target) of the following
        call: its copy runs under this step's compute.  `host_loss` = pinned fp32 tensor of one element that receives this step's
        loss asynchronously (valid after a synchronize; `join=True` makes the current stream wait for it)."""
        import ctypes as C
        B = host_input.shape[0]
        if host_input.is_cuda or host_target.is_cuda or host_input.dtype != torch.float32 or host_target.dtype != torch.float32:
            raise RuntimeError("training_step_host takes fp32 host tensors")
        need = int(self.lib.tcnn_b200_host_staging_bytes(self.native.handle, B))
        st = getattr(self, "_host_staging", None)
        if st is None or st.numel() < need:
            st = self._host_staging = torch.empty(need, dtype=torch.uint8, device=self.device)
            self._host_views = {}
        dx, dt, dl = C.c_void_p(), C.c_void_p(), C.c_void_p()
        nx, nt = (next_batch if next_batch is not None else (None, None))
        check(self.lib.tcnn_b200_host_batch_begin(self.native.handle, _stream_ptr(), B, host_input.data_ptr(), host_target.data_ptr(),
                                                  nx.data_ptr() if nx is not None else None, nt.data_ptr() if nt is not None else None,
                                                  st.data_ptr(), C.byref(dx), C.byref(dt), C.byref(dl)))
        key = (dx.value, B)
        views = self._host_views.get(key)
        if views is None:
            base = st.data_ptr()
            def view(p, n, shape):
                return st[p - base: p - base + 4 * n].view(torch.float32).view(shape)
            views = self._host_views[key] = (view(dx.value, B * host_input.shape[1], (B, host_input.shape[1])),
                                             view(dt.value, B * host_target.shape[1], (B, host_target.shape[1])), view(dl.value, 1, (1,)))
        keep = self.loss_value
        self.loss_value = views[2]
        try:
            self.training_step(views[0], views[1], **step_kwargs)
        finally:
            self.loss_value = keep
            check(self.lib.tcnn_b200_host_batch_end(self.native.handle, _stream_ptr(), host_loss.data_ptr() if host_loss is not None else None, 1 if join else 0))
        return self

    # static CUDA graphs behind the fused step (cuda_graph.h:65-153 of the reference): 1 on, 0 off, -1 environment TCNN_B200_GRAPH
    def set_graph_mode(self, mode: int):
        check(self.lib.tcnn_b200_plan_set_graph_mode(self.native.handle, int(mode)))

    def graph_stats(self) -> tuple[int, int]:
        import ctypes as C
        captures, replays = C.c_uint32(0), C.c_uint32(0)
        check(self.lib.tcnn_b200_plan_graph_stats(self.native.handle, C.byref(captures), C.byref(replays)))
        return captures.value, replays.value

    def update_hyperparams(self, params: dict):
        check(self.lib.tcnn_b200_plan_update_optimizer(self.native.handle, json.dumps(params).encode()))
        self.native.refresh_info()

    # Network::inference (object.h:147-176): fp32 [B, O] output
    def inference(self, input: torch.Tensor, output: torch.Tensor | None = None) -> torch.Tensor:
        B = input.shape[0]
        if output is None:
            output = torch.empty(B, int(self.info.n_output_dims), dtype=torch.float32, device=self.device)
        ws = self.native.workspace(B, self.device)
        check(self.lib.tcnn_b200_inference(self.native.handle, _stream_ptr(), B, ptr(input), ptr(self.params), None, ptr(output), ptr(ws)))
        return output


@dataclass
class TrainableModel:
    """config.h:46-51: {loss, optimizer, network, trainer}; loss and optimizer live inside the plan here."""
    network: NativeModule
    trainer: Trainer


def create_from_config(n_input_dims: int, n_output_dims: int, config: dict, seed: int = 1337, device="cuda", process_group=None, **trainer_kwargs) -> TrainableModel:
    """config.h:53-63.  config = {"loss": {...}, "optimizer": {...}, "encoding": {...}, "network": {...}}."""
    with torch.cuda.device(device):
        native = NativeModule(n_input_dims, n_output_dims, config)
        return TrainableModel(network=native, trainer=Trainer(native, seed=seed, device=device, process_group=process_group, **trainer_kwargs))
