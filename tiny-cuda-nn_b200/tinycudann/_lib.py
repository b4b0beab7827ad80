"""ctypes binding of libtcnn_b200.so (include/tcnn_b200.h).  This is the stub a maintainer of the reference's
bindings/torch would add in place of the pybind11 extension `_C` (bindings/torch/tinycudann/bindings.cpp).

There is no fallback: if the shared library is missing or a call fails, a RuntimeError is raised."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TCNN_B200_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libtcnn_b200.so")  # env: A/B builds


class Info(C.Structure):
    _fields_ = [
        ("n_input_dims", C.c_uint32),
        ("n_output_dims", C.c_uint32),
        ("padded_output_width", C.c_uint32),
        ("n_levels", C.c_uint32),
        ("n_features_per_level", C.c_uint32),
        ("log2_hashmap_size", C.c_uint32),
        ("base_resolution", C.c_uint32),
        ("per_level_scale", C.c_float),
        ("encoded_width", C.c_uint32),
        ("n_neurons", C.c_uint32),
        ("n_hidden_layers", C.c_uint32),
        ("n_params", C.c_uint64),
        ("n_params_mlp", C.c_uint64),
        ("n_params_grid", C.c_uint64),
        ("loss_type", C.c_int),
        ("learning_rate", C.c_float),
        ("beta1", C.c_float),
        ("beta2", C.c_float),
        ("epsilon", C.c_float),
        ("l2_reg", C.c_float),
        ("relative_decay", C.c_float),
        ("absolute_decay", C.c_float),
        ("level_offset", C.c_uint32 * 33),
        ("level_scale", C.c_float * 32),
        ("level_resolution", C.c_uint32 * 32),
    ]


_P = C.c_void_p


class DpComm(C.Structure):
    """tcnn_b200_dp_comm (include/tcnn_b200.h)."""
    _fields_ = [
        ("rank", C.c_uint32), ("world_size", C.c_uint32), ("epoch", C.c_uint32), ("reserved", C.c_uint32),
        ("peer_grads", C.c_void_p * 8), ("peer_params", C.c_void_p * 8), ("peer_flags", C.c_void_p * 8),
        ("grads_multicast", C.c_void_p), ("params_multicast", C.c_void_p),
    ]


_SIGNATURES = {
    # name: (restype, argtypes)
    "tcnn_b200_last_error": (C.c_char_p, []),
    "tcnn_b200_abi_version": (C.c_uint32, []),
    "tcnn_b200_arch": (C.c_char_p, []),
    "tcnn_b200_batch_size_granularity": (C.c_uint32, []),
    "tcnn_b200_plan_create": (C.c_int, [C.c_char_p, C.c_uint32, C.c_uint32, C.POINTER(_P)]),
    "tcnn_b200_plan_destroy": (None, [_P]),
    "tcnn_b200_plan_info": (C.c_int, [_P, C.POINTER(Info)]),
    "tcnn_b200_plan_set_max_level": (C.c_int, [_P, C.c_float]),
    "tcnn_b200_plan_update_optimizer": (C.c_int, [_P, C.c_char_p]),
    "tcnn_b200_workspace_bytes": (C.c_size_t, [_P, C.c_uint32]),
    "tcnn_b200_init_params": (C.c_int, [_P, _P, C.c_uint64, C.c_int, _P]),
    "tcnn_b200_cast_params": (C.c_int, [_P, C.c_uint64, _P, _P]),
    "tcnn_b200_generate_random_uniform": (C.c_int, [_P, C.c_uint64, C.c_uint64, C.c_uint64, _P, C.c_float, C.c_float]),
    "tcnn_b200_encode_forward": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P]),
    "tcnn_b200_encode_backward": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P, C.c_int, _P]),
    "tcnn_b200_encode_backward_noalloc": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P, C.c_int]),
    "tcnn_b200_grid_indices": (C.c_int, [_P, _P, C.c_uint32, _P, _P]),
    "tcnn_b200_mlp_forward": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P]),
    "tcnn_b200_mlp_backward": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P, _P, _P, _P]),
    "tcnn_b200_inference": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P, _P, _P]),
    "tcnn_b200_forward": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P, _P]),
    "tcnn_b200_backward": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P, _P, _P, _P]),
    "tcnn_b200_backward_backward_input": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P, _P, _P, _P, _P, _P]),
    "tcnn_b200_training_fwd_bwd": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P, _P, C.c_float, C.c_int, _P, _P, _P]),
    "tcnn_b200_adam_step": (C.c_int, [_P, _P, C.c_float, C.c_uint32, _P, _P, _P, _P, _P, _P, C.c_int]),
    "tcnn_b200_adam_step_range": (C.c_int, [_P, _P, C.c_float, C.c_uint32, _P, _P, _P, _P, _P, _P, C.c_int, C.c_uint64, C.c_uint64]),
    "tcnn_b200_dp_shard": (C.c_int, [_P, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "tcnn_b200_dp_flag_bytes": (C.c_uint32, []),
    "tcnn_b200_dp_flag_error_word": (C.c_uint32, []),
    "tcnn_b200_training_step_dp": (C.c_int, [_P, _P, C.c_uint32, _P, _P, C.c_float, C.c_uint32, _P, _P, _P, _P, _P, _P, _P, C.POINTER(DpComm)]),
    "tcnn_b200_training_step": (C.c_int, [_P, _P, C.c_uint32, _P, _P, C.c_float, C.c_uint32, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "tcnn_b200_plan_set_graph_mode": (C.c_int, [_P, C.c_int]),
    "tcnn_b200_plan_graph_stats": (C.c_int, [_P, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "tcnn_b200_host_staging_bytes": (C.c_size_t, [_P, C.c_uint32]),
    "tcnn_b200_host_batch_begin": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P, _P, _P, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "tcnn_b200_host_batch_end": (C.c_int, [_P, _P, _P, C.c_int]),
    "tcnn_b200_backward_mode": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P, _P, _P, _P, C.c_int]),
    "tcnn_b200_plan_set_max_level_gpu": (C.c_int, [_P, _P]),
    "tcnn_b200_profile_enable": (C.c_int, [_P, C.c_int]),
    "tcnn_b200_profile_read": (C.c_int, [_P, C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.c_uint32]),
}

_lib = None


def exported_symbols():
    return sorted(_SIGNATURES)


def lib():
    """Load (once) and return the shared library; raises if it has not been built (see __graft_entry__.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not found: build it with `python __graft_entry__.py build` (no CPU fallback exists)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc: int):
    if rc != 0:
        raise RuntimeError(lib().tcnn_b200_last_error().decode("utf-8", "replace"))


def ptr(t):
    """Device (or None) pointer of a torch tensor as void*."""
    return None if t is None else C.c_void_p(t.data_ptr())
