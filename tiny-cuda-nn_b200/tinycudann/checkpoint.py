"""Snapshot format of the reference's Trainer (include/tiny-cuda-nn/trainer.h:275-315) and AdamOptimizer
(optimizers/adam.h:278-299): a JSON object whose arrays are binary blobs (gpu_memory_json.h:36-72),

    {"n_params": P, "params_type": "__half" | "float", "params_binary": <bytes>,
     "optimizer": {"current_step", "base_learning_rate", "first_moments_binary", "second_moments_binary", "param_steps_binary"}}

Hosts of the reference write it with nlohmann `json::to_msgpack`; `msgpack.packb(snapshot)` produces the same bytes layout
(map of str -> int / str / bin), and `msgpack.unpackb` of a reference file gives a dict this module reads.  Blobs may also
come as {"bytes": [...]} (nlohmann's JSON text form of a binary value), which from_json accepts too."""
from __future__ import annotations

import numpy as np

_DTYPES = {"__half": np.float16, "float": np.float32}


def blob(a: np.ndarray) -> bytes:
    return np.ascontiguousarray(a).tobytes()


def unblob(b, dtype) -> np.ndarray:
    if isinstance(b, dict):  # {"bytes": [...], "subtype": null}
        b = bytes(bytearray(b["bytes"]))
    if not isinstance(b, (bytes, bytearray, memoryview)):
        raise RuntimeError("Invalid json type: must be either binary or object")
    return np.frombuffer(bytes(b), dtype=dtype).copy()


def pack(params_fp16: np.ndarray, optimizer: dict | None = None) -> dict:
    data = {"n_params": int(params_fp16.size), "params_type": "__half", "params_binary": blob(params_fp16.astype(np.float16))}
    if optimizer is not None:
        data["optimizer"] = {
            "current_step": int(optimizer["current_step"]),
            "base_learning_rate": float(optimizer["base_learning_rate"]),
            "first_moments_binary": blob(optimizer["first_moments"].astype(np.float32)),
            "second_moments_binary": blob(optimizer["second_moments"].astype(np.float32)),
            "param_steps_binary": blob(optimizer["param_steps"].astype(np.uint32)),
        }
    return data


def unpack(data: dict):
    """-> (params ndarray (fp16 or fp32, as stored), optimizer dict or None)"""
    ptype = data.get("params_type", "__half")
    if ptype not in _DTYPES:
        raise RuntimeError("Trainer: snapshot parameters must be of type float of __half")
    params = unblob(data["params_binary"], _DTYPES[ptype])
    opt = None
    if "optimizer" in data:
        o = data["optimizer"]
        second = unblob(o["second_moments_binary"], np.float32)
        opt = {
            "current_step": int(o["current_step"]),
            "base_learning_rate": float(o["base_learning_rate"]),
            "first_moments": unblob(o["first_moments_binary"], np.float32),
            "second_moments": second,
            "param_steps": unblob(o["param_steps_binary"], np.uint32) if "param_steps_binary" in o else np.zeros(second.size, np.uint32),
        }
    return params, opt
