"""PyTorch module surface of the reference's bindings (bindings/torch/tinycudann/modules.py:162-329), on top of
libtcnn_b200.so.  Same constructor arguments, same `.params` / `.n_output_dims` / `.dtype` / `.loss_scale`
attributes, same batch padding to 256 and output slicing (modules.py:176-192), same loss-scaling convention in the
autograd function (modules.py:91-130).  PyTorch is used for device memory and streams only; all compute goes
through the C ABI."""
from __future__ import annotations

import json
import warnings
import weakref

import torch

from . import _lib
from ._lib import check, ptr


def _stream_ptr():
    return _lib.C.c_void_p(torch.cuda.current_stream().cuda_stream)


_live_modules = weakref.WeakSet()  # for free_temporary_memory(): every NativeModule that may hold a cached workspace


class NativeModule:
    """Plays the role of the pybind11 `Module` (bindings/torch/tinycudann/bindings.cpp:60-262) over a C-ABI plan."""

    def __init__(self, n_input_dims: int, n_output_dims: int, config: dict):
        self.lib = _lib.lib()
        self.config = config
        handle = _lib.C.c_void_p()
        check(self.lib.tcnn_b200_plan_create(json.dumps(config).encode(), n_input_dims, n_output_dims, _lib.C.byref(handle)))
        self.handle = handle
        # Precision of parameters and output (cpp_api.h:71-75): fp32 for a stand-alone encoding created with "precision": "fp32"
        self.param_dtype = torch.float32 if str(config.get("precision", "fp16")).lower() == "fp32" else torch.half
        self.info = _lib.Info()
        check(self.lib.tcnn_b200_plan_info(self.handle, _lib.C.byref(self.info)))
        self._ws = {}
        _live_modules.add(self)

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self.lib.tcnn_b200_plan_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    def set_max_level(self, max_level: float):
        """GridEncoding::set_max_level (grid_interface.h:101-107): levels >= max_level * n_levels encode to zero."""
        check(self.lib.tcnn_b200_plan_set_max_level(self.handle, float(max_level)))

    def set_max_level_gpu(self, max_level_per_sample: torch.Tensor | None):
        """GridEncoding::set_max_level_gpu (grid_interface.h:113-115): one fp32 max level per sample of the batches to come, on the
        device; the tensor is kept alive here because the library keeps the pointer."""
        if max_level_per_sample is not None and (not max_level_per_sample.is_cuda or max_level_per_sample.dtype != torch.float32):
            raise RuntimeError("set_max_level_gpu takes a float32 CUDA tensor")
        self._max_level_gpu = max_level_per_sample
        check(self.lib.tcnn_b200_plan_set_max_level_gpu(self.handle, ptr(max_level_per_sample)))

    def refresh_info(self):
        check(self.lib.tcnn_b200_plan_info(self.handle, _lib.C.byref(self.info)))

    # --- queries (cpp_api.h:100-115) ---
    def n_params(self) -> int:
        return int(self.info.n_params)

    def n_output_dims(self) -> int:
        return int(self.info.padded_output_width)

    def hyperparams(self) -> dict:
        return self.config

    def workspace(self, batch_size: int, device, fresh: bool = False) -> torch.Tensor:
        n = int(self.lib.tcnn_b200_workspace_bytes(self.handle, batch_size))
        if fresh:
            return torch.empty(n, dtype=torch.uint8, device=device)
        # ONE grow-only scratch buffer per (device, stream), like the reference's per-stream arena (gpu_memory.h:699-720): callers with
        # varying batch sizes (NeRF ray batches) re-use it instead of piling up one buffer per size; free_temporary_memory() drops it
        key = (str(device), torch.cuda.current_stream(device).cuda_stream if torch.device(device).type == "cuda" else 0)
        ws = self._ws.get(key)
        if ws is None or ws.numel() < n:
            self._ws[key] = ws = torch.empty(n, dtype=torch.uint8, device=device)
        return ws

    def free_workspaces(self):
        self._ws.clear()

    def initial_params(self, seed: int, device="cuda", trainer_seeding: bool = False) -> torch.Tensor:
        p = torch.empty(self.n_params(), dtype=torch.float32, device=device)
        check(self.lib.tcnn_b200_init_params(self.handle, _stream_ptr(), seed, 1 if trainer_seeding else 0, ptr(p)))
        return p

    # --- compute ---
    def fwd(self, x: torch.Tensor, params: torch.Tensor, need_ctx: bool):
        B = x.shape[0]
        out = torch.empty(B, self.n_output_dims(), dtype=self.param_dtype, device=x.device)
        if need_ctx:
            ws = self.workspace(B, x.device, fresh=True)  # owned by the autograd context, like the reference's Context
            check(self.lib.tcnn_b200_forward(self.handle, _stream_ptr(), B, ptr(x), ptr(params), ptr(out), ptr(ws)))
            return ws, out
        ws = self.workspace(B, x.device)
        check(self.lib.tcnn_b200_inference(self.handle, _stream_ptr(), B, ptr(x), ptr(params), ptr(out), None, ptr(ws)))
        return None, out

    def bwd(self, ws, x, params, doutput, want_params_grad: bool, want_input_grad: bool = False):
        """-> (dL_dinput fp32 [B, n_input_dims] or None, dL_dparams [n_params] in the parameter precision or None)"""
        B = x.shape[0]
        g = torch.empty(self.n_params(), dtype=self.param_dtype, device=x.device) if want_params_grad else None
        dx = torch.empty(B, int(self.info.n_input_dims), dtype=torch.float32, device=x.device) if want_input_grad else None
        if g is not None or dx is not None:
            check(self.lib.tcnn_b200_backward(self.handle, _stream_ptr(), B, ptr(x), ptr(params), ptr(doutput), ptr(g), ptr(dx), ptr(ws)))
        return dx, g


    def bwd_bwd_input(self, x, params, dL_ddLdinput, doutput, want_doutput_grad=True, want_params_grad=True, want_input_grad=True):
        """DifferentiableObject::backward_backward_input (object.h:208-240): -> (dL_ddLdoutput fp16 [B, width], dL_dparams fp16, dL_dinput fp32)."""
        B = x.shape[0]
        dd = torch.empty(B, self.n_output_dims(), dtype=torch.half, device=x.device) if want_doutput_grad else None
        g = torch.empty(self.n_params(), dtype=torch.half, device=x.device) if want_params_grad else None
        dx = torch.empty(B, int(self.info.n_input_dims), dtype=torch.float32, device=x.device) if want_input_grad else None
        check(self.lib.tcnn_b200_backward_backward_input(self.handle, _stream_ptr(), B, ptr(x), ptr(params), ptr(dL_ddLdinput), ptr(doutput),
                                                         ptr(dd), ptr(g), ptr(dx), None))
        return dd, g, dx


def _empty_like(t):
    return torch.empty([], dtype=t.dtype, device=t.device)


def _none_if_empty(t):
    return None if t.dim() == 0 else t


class _module_function(torch.autograd.Function):
    @staticmethod
    def forward(ctx, native, x, params, loss_scale, need_ctx):
        ctx.set_materialize_grads(False)
        ws, output = native.fwd(x, params, need_ctx)
        ctx.save_for_backward(x, params)
        ctx.native, ctx.ws, ctx.loss_scale = native, ws, loss_scale
        return output

    @staticmethod
    def backward(ctx, doutput):
        if doutput is None:
            return None, None, None, None, None
        x, params = ctx.saved_tensors
        # the backward pass is itself a differentiable function (second order of the grid encoding, modules.py:120-160 of the reference)
        dx, g = _module_function_backward.apply(ctx, doutput, x, params)
        return None, _none_if_empty(dx), _none_if_empty(g), None, None


class _module_function_backward(torch.autograd.Function):
    """dL/dinput and dL/dparams from dL/doutput.  Its own backward supports, like the reference, the gradients that flow through
    dL/dinput: d(dL_dinput)/d(dL_doutput), d(dL_dinput)/d(params), d(dL_dinput)/d(input); nothing through dL/dparams."""

    @staticmethod
    def forward(ctx, fwd, doutput, x, params):
        ctx.fwd = fwd
        ctx.save_for_backward(x, params, doutput)
        with torch.no_grad():
            scaled = (doutput.to(torch.float) * fwd.loss_scale).to(fwd.native.param_dtype).contiguous()
            dx, g = fwd.native.bwd(fwd.ws, x, params, scaled, fwd.needs_input_grad[2], fwd.needs_input_grad[1])
            dx = _empty_like(x) if dx is None else dx / fwd.loss_scale
            g = _empty_like(params) if g is None else g / fwd.loss_scale
        return dx, g

    @staticmethod
    def backward(ctx, ddx, dg):
        if ddx is None or ddx.dim() == 0:
            return None, None, None, None
        x, params, doutput = ctx.saved_tensors
        fwd = ctx.fwd
        with torch.no_grad():
            scaled = (doutput.to(torch.float) * fwd.loss_scale).to(torch.half).contiguous()
            d_doutput, d_params, d_x = fwd.native.bwd_bwd_input(x, params, ddx.to(torch.float).contiguous(), scaled,
                                                                 ctx.needs_input_grad[1], ctx.needs_input_grad[3], ctx.needs_input_grad[2])
            # d_doutput depends on ddx only; the other two on ddx * doutput, which carried the loss scale
            d_params = None if d_params is None else d_params / fwd.loss_scale
            d_x = None if d_x is None else d_x / fwd.loss_scale
        return None, d_doutput, d_x, d_params


class Module(torch.nn.Module):
    def __init__(self, seed=1337):
        super().__init__()
        self.native_tcnn_module = self._native_tcnn_module()
        self.dtype = self.native_tcnn_module.param_dtype
        self.seed = seed
        initial_params = self.native_tcnn_module.initial_params(seed)
        self.params = torch.nn.Parameter(initial_params, requires_grad=True)
        self.loss_scale = 128.0 if self.dtype == torch.half else 1.0  # default_loss_scale(precision), cpp_api.h:76 / common.h:229-233

    def forward(self, x):
        if not x.is_cuda:
            warnings.warn("input must be a CUDA tensor, but isn't. This indicates suboptimal performance.")
            x = x.cuda()
        batch_size = x.shape[0]
        g = int(_lib.lib().tcnn_b200_batch_size_granularity())
        padded = (batch_size + g - 1) // g * g
        x_padded = x if batch_size == padded else torch.nn.functional.pad(x, [0, 0, 0, padded - batch_size])
        output = _module_function.apply(
            self.native_tcnn_module,
            x_padded.to(torch.float).contiguous(),
            self.params.to(self.dtype).contiguous(),
            self.loss_scale,
            torch.is_grad_enabled() and (self.params.requires_grad or x.requires_grad),  # grad mode is off inside Function.forward
        )
        return output[:batch_size, : self.n_output_dims]

    def __getstate__(self):
        state = self.__dict__.copy()
        del state["native_tcnn_module"]
        return state

    def __setstate__(self, state):
        self.__dict__.update(state)
        self.native_tcnn_module = self._native_tcnn_module()

    def extra_repr(self):
        return f"n_input_dims={self.n_input_dims}, n_output_dims={self.n_output_dims}, seed={self.seed}, dtype={self.dtype}, hyperparams={self.native_tcnn_module.hyperparams()}"


class NetworkWithInputEncoding(Module):
    """Input encoding followed by a neural network (modules.py:209-247 of the reference).  Supported here:
    HashGrid encoding + FullyFusedMLP (64 neurons, ReLU, linear output)."""

    def __init__(self, n_input_dims, n_output_dims, encoding_config, network_config, seed=1337):
        self.n_input_dims = n_input_dims
        self.n_output_dims = n_output_dims
        self.encoding_config = encoding_config
        self.network_config = network_config
        super().__init__(seed=seed)

    def _native_tcnn_module(self):
        return NativeModule(self.n_input_dims, self.n_output_dims, {"encoding": self.encoding_config, "network": self.network_config})


class Network(Module):
    """Neural network (modules.py:251-283 of the reference; src/cpp_api.cu:151-153: Identity encoding + network).  Supported
    here: FullyFusedMLP with 64 neurons, 1-4 hidden layers, ReLU, linear output, n_input_dims <= 64, n_output_dims <= 16.
    Gradients w.r.t. the input are available (identity_backward)."""

    def __init__(self, n_input_dims, n_output_dims, network_config, seed=1337):
        self.n_input_dims = n_input_dims
        self.n_output_dims = n_output_dims
        self.network_config = network_config
        super().__init__(seed=seed)

    def _native_tcnn_module(self):
        return NativeModule(self.n_input_dims, self.n_output_dims, {"network": self.network_config})


class Encoding(Module):
    """Input encoding (modules.py:285-329 of the reference; src/cpp_api.cu:156-165: no padding of the output).  Supported
    here: grid encodings; dtype None / torch.half (fp16 parameters and output) or torch.float32 (fp32 throughout; first order)."""

    def __init__(self, n_input_dims, encoding_config, seed=1337, dtype=None):
        self.n_input_dims = n_input_dims
        self.encoding_config = encoding_config
        if dtype not in (None, torch.float16, torch.float32):
            raise ValueError(f"Encoding only supports fp32 or fp16 precision, but got {dtype}")
        self.precision = "fp32" if dtype == torch.float32 else "fp16"
        super().__init__(seed=seed)
        self.n_output_dims = int(self.native_tcnn_module.info.n_output_dims)

    def _native_tcnn_module(self):
        return NativeModule(self.n_input_dims, 0, {"encoding": self.encoding_config, "precision": self.precision})


def free_temporary_memory():
    """tcnn.free_temporary_memory (modules.py:77-81 -> free_all_gpu_memory_arenas): the C ABI owns no temporaries; the scratch buffers are
    torch tensors cached per module and stream (NativeModule.workspace) and are released here."""
    import gc

    for m in list(_live_modules):
        m.free_workspaces()
    gc.collect()
    torch.cuda.empty_cache()
