"""ctypes binding of ``libptdb200.so`` (declarations mirror ``include/ptdb200.h``).

The library is plain C ABI: no torch C++ ABI coupling, compiles in seconds, and the same ``.so``
serves Python, tests and foreign callers.  ``call(name, *args)`` accepts tensors (passed as device
pointers), ``None`` (NULL), ints and floats; a non-zero return raises ``RuntimeError`` with the
library's message.  On a box with a GPU a missing library is a hard error - there is no silent
fallback to eager PyTorch for the CUDA ops.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_longlong, c_void_p
from pathlib import Path
from typing import Any, Optional

import torch

P, I, L, F = c_void_p, c_int, c_longlong, c_float

_SIGS: dict[str, list] = {
    "ptd_version": [],
    "ptd_device_sm_count": [],
    "ptd_set_pdl_mode": [I],
    "ptd_conv_fprop": [P, P, P, P, P, I, I, I, I, I, I, I, I, I, P],
    "ptd_conv_dgrad": [P, P, P, P, I, I, I, I, I, I, I, I, I, P],
    "ptd_conv_dgrad_accum": [P, P, P, I, I, I, I, I, I, I, I, I, P],
    "ptd_conv_dgrad_fused": [P, P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, P],
    "ptd_conv_wgrad": [P, P, P, I, I, I, I, I, I, I, I, I, P],
    "ptd_bottleneck_fwd": [P, P],
    "ptd_bottleneck_bwd": [P, P],
    "ptd_stem_fprop": [P, P, P, P, P, I, I, I, P],
    "ptd_stem_wgrad": [P, P, P, I, I, I, P],
    "ptd_stem_s2d_fprop": [P, P, P, P, P, I, I, I, P],
    "ptd_stem_s2d_wgrad": [P, P, P, I, I, I, P],
    "ptd_linear_fwd": [P, P, P, P, I, I, I, P],
    "ptd_linear_dgrad": [P, P, P, I, I, I, P],
    "ptd_linear_wgrad": [P, P, P, I, I, I, P],
    "ptd_bn_stats": [P, P, P, L, I, P],
    "ptd_bn_apply": [P, P, P, P, P, P, P, P, P, P, P, L, I, F, F, I, P],
    "ptd_bn_apply_mask": [P, P, P, P, P, P, P, P, P, P, P, P, L, I, F, F, I, P],
    "ptd_bn_apply_dual": [P] * 20 + [L, I, F, F, P],
    "ptd_bn_bwd_dual": [P] * 14 + [L, I, P],
    "ptd_conv_dgrad_masked_addend": [P, P, P, P, P, I, I, I, I, I, P],
    "ptd_bn_relu_maxpool_fwd": [P, P, P, P, P, P, P, P, P, P, P, I, I, I, I, F, F, P],
    "ptd_bn_relu_maxpool_bwd": [P, P, P, P, P, P, P, P, I, I, I, I, P],
    "ptd_affine_act": [P, P, P, P, P, L, I, I, P],
    "ptd_bn_bwd_reduce": [P, P, P, P, P, P, P, L, I, I, P],
    "ptd_bn_bwd_apply": [P, P, P, P, P, P, P, P, P, P, L, I, I, P],
    "ptd_bn_backward_static": [P, P, P, P, P, P, P, P, P, P, L, I, I, P],
    "ptd_maxpool3x3s2_fwd": [P, P, P, I, I, I, I, P],
    "ptd_maxpool3x3s2_bwd": [P, P, P, I, I, I, I, P],
    "ptd_avgpool_fwd": [P, P, I, I, I, P],
    "ptd_avgpool_bwd": [P, P, I, I, I, P],
    "ptd_nchw3_f32_to_nhwc4_bf16": [P, P, I, I, I, P],
    "ptd_nchw_f32_to_nhwc_bf16": [P, P, I, I, I, P],
    "ptd_nhwc_bf16_to_nchw_f32": [P, P, I, I, I, P],
    "ptd_cast_f32_to_bf16": [P, P, L, P],
    "ptd_stem_pack_weight": [P, P, L, L, L, L, P],
    "ptd_nchw3_f32_to_s2d_bf16": [P, P, I, I, I, P],
    "ptd_stem_pack_weight_s2d": [P, P, L, L, L, L, P],
    "ptd_stem_unpack_wgrad_s2d": [P, P, L, L, L, L, I, P],
    "ptd_stem_unpack_wgrad": [P, P, L, L, L, L, I, P],
    "ptd_softmax_ce": [P, P, P, P, P, I, I, I, I, F, P],
    "ptd_multi_sgd": [P, P, I, F, F, F, F, I, P],
    "ptd_multi_adam": [P, P, I, F, F, F, F, F, I, F, P],
    "ptd_multi_sgd_dev": [P, P, I, P, F, F, F, I, P],
    "ptd_multi_adam_dev": [P, P, I, P, F, F, F, F, P, F, P],
    "ptd_multi_cast_bf16": [P, P, I, P],
    "ptd_multi_zero_grad": [P, P, I, P],
    "ptd_opt_chunk_elems": [],
    "ptd_small_linear_fwd": [P, P, P, P, I, I, I, I, P],
    "ptd_small_linear_bwd": [P, P, P, P, P, P, I, I, I, P],
    "ptd_mlp_fwd": [P, P, P, P, P, I, I, P],
}



class BottleneckFwd(ctypes.Structure):
    """Mirror of PtdBottleneckFwd (include/ptdb200.h)."""
    _fields_ = [("N", I), ("H", I), ("W", I), ("inplanes", I), ("planes", I), ("stride", I), ("has_ds", I),
                ("eps", F), ("momentum", F), ("x", P), ("w", P * 4), ("gamma", P * 4), ("beta", P * 4),
                ("running_mean", P * 4), ("running_var", P * 4), ("y", P * 4), ("z", P * 4), ("stats", P), ("saved", P), ("mask", P * 4)]


class BottleneckBwd(ctypes.Structure):
    """Mirror of PtdBottleneckBwd (include/ptdb200.h)."""
    _fields_ = [("N", I), ("H", I), ("W", I), ("inplanes", I), ("planes", I), ("stride", I), ("has_ds", I),
                ("x", P), ("dz", P), ("w", P * 4), ("gamma", P * 4), ("y", P * 4), ("z", P * 4), ("saved", P),
                ("sums", P), ("dw_all", P), ("dw_all_bytes", L), ("dw", P * 4), ("dy", P * 4), ("g_id", P), ("dz2", P),
                ("dz1", P), ("dx", P), ("mask", P * 4), ("gdst", P * 8)]


class SgdEpilogue(ctypes.Structure):
    """Mirror of PtdSgdEpilogue (include/ptdb200.h)."""
    _fields_ = [("param", P), ("momentum_buf", P), ("shadow_bf16", P), ("lr_dev", P), ("momentum", F), ("weight_decay", F)]


class BucketComm(ctypes.Structure):
    """Mirror of PtdBucketComm (include/ptdb200.h)."""
    _fields_ = [("algo", I), ("rank", I), ("world", I), ("blocks", I), ("numel", L), ("scale", F), ("local", P),
                ("bufs", P * 8), ("mc_buf", P), ("stages", P * 8), ("mc_stage", P), ("pads", P * 8), ("mc_pad", P),
                ("fuse_sgd", I), ("sgd", SgdEpilogue)]


_LIB: Optional[ctypes.CDLL] = None
_LOAD_ERROR: Optional[str] = None
LIB_PATH = Path(__file__).resolve().parent / "lib" / "libptdb200.so"


def _load() -> Optional[ctypes.CDLL]:
    global _LIB, _LOAD_ERROR
    if _LIB is not None or _LOAD_ERROR is not None:
        return _LIB
    path = Path(os.environ.get("PTDB200_LIB", LIB_PATH))
    if not path.exists():
        try:
            from ._build import build
            build()
        except Exception as e:  # noqa: BLE001 - surfaced through available()/require()
            _LOAD_ERROR = f"{path} missing and in-tree build failed: {e}"
            return None
    try:
        lib = ctypes.CDLL(str(path))
    except OSError as e:
        _LOAD_ERROR = f"cannot load {path}: {e}"
        return None
    lib.ptd_last_error.restype = c_char_p
    lib.ptd_last_error.argtypes = []
    for name, sig in _SIGS.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.restype = c_int
        fn.argtypes = sig
    _LIB = lib
    return _LIB


def available() -> bool:
    return _load() is not None


def load_error() -> Optional[str]:
    _load()
    return _LOAD_ERROR


def require() -> ctypes.CDLL:
    lib = _load()
    if lib is None:
        raise RuntimeError(f"pytorchdistributed_b200 native library unavailable: {_LOAD_ERROR}")
    return lib


def register(name: str, sig: list) -> None:
    """Declare an entry point added by another module (e.g. the peer-to-peer runtime)."""
    _SIGS[name] = sig
    if _LIB is not None and hasattr(_LIB, name):
        fn = getattr(_LIB, name)
        fn.restype = c_int
        fn.argtypes = sig


def _conv(a: Any):
    if a is None:
        return None
    if isinstance(a, torch.Tensor):
        return a.data_ptr()
    if isinstance(a, ctypes.Structure):
        return ctypes.addressof(a)
    return a


def stream_ptr(device: Optional[torch.device] = None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


LAUNCHES = 0   # entry-point calls that launch device work (bench.py reports it as gpu_launches)


_PEER_OK = ("ptd_peer_copy", "ptd_peer_copy_bulk", "ptd_reduce_sum_f32", "ptd_allreduce", "ptd_broadcast", "ptd_shard",
            "ptd_p2p", "ptd_vmm")      # entry points whose tensors legitimately live on other GPUs


def call(name: str, *args: Any) -> None:
    global LAUNCHES
    lib = require()
    LAUNCHES += 1
    if not name.startswith(_PEER_OK):
        for a in args:
            if isinstance(a, torch.Tensor):
                if a.is_cuda and a.device.index != torch.cuda.current_device():
                    raise RuntimeError(f"{name}: tensor on {a.device} but the current device is cuda:"
                                       f"{torch.cuda.current_device()} - kernels launch on the current device "
                                       "(wrap the call in torch.cuda.device(...))")
                break
    rc = getattr(lib, name)(*[_conv(a) for a in args])
    if rc != 0:
        msg = lib.ptd_last_error()
        raise RuntimeError(f"{name} failed ({rc}): {msg.decode() if msg else '?'}")


def call_ret(name: str, *args: Any) -> int:
    """Call an entry point whose return value is data (not an error code)."""
    lib = require()
    return getattr(lib, name)(*[_conv(a) for a in args])
