"""NCCL point-to-point between the GPUs of ONE process (model-parallel stage boundary).

``torch.distributed`` needs one process per rank; the reference's split model lives in a single
process with one stage per device (``R/03 模型并行/03_model_parallel.ipynb#cell5``: ``.to('cuda:1')``
peer copies on the compute stream).  Here the activation / gradient hand-off is an
``ncclSend``/``ncclRecv`` pair on per-device side streams, through communicators created with
``ncclCommInitAll`` on the NCCL library torch already loaded (ctypes; no extra build step).
"""
from __future__ import annotations

import ctypes
import ctypes.util
import glob
import os
from typing import Dict, List, Optional, Sequence

import torch

_NCCL_UINT8 = 1
_lib: Optional[ctypes.CDLL] = None


def _load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    cands = ["libnccl.so.2"]
    site = os.path.dirname(os.path.dirname(torch.__file__))
    cands += glob.glob(os.path.join(site, "nvidia", "nccl", "lib", "libnccl.so*"))
    err = None
    for c in cands:
        try:
            lib = ctypes.CDLL(c)
            break
        except OSError as e:
            err = e
    else:
        raise RuntimeError(f"cannot load NCCL: {err}")
    lib.ncclCommInitAll.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    lib.ncclSend.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    lib.ncclRecv.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    lib.ncclCommDestroy.argtypes = [ctypes.c_void_p]
    lib.ncclGetErrorString.restype = ctypes.c_char_p
    lib.ncclGetErrorString.argtypes = [ctypes.c_int]
    _lib = lib
    return lib


def _check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed: {_load().ncclGetErrorString(rc).decode()}")


class LocalNcclGroup:
    """One communicator per device of this process; ``send_recv`` moves a tensor between two of them."""

    def __init__(self, devices: Sequence[int]) -> None:
        lib = _load()
        self.devices = list(devices)
        n = len(self.devices)
        comms = (ctypes.c_void_p * n)()
        devs = (ctypes.c_int * n)(*self.devices)
        _check(lib.ncclCommInitAll(ctypes.cast(comms, ctypes.c_void_p), n, ctypes.cast(devs, ctypes.c_void_p)),
               "ncclCommInitAll")
        self.comms: Dict[int, int] = {d: comms[i] for i, d in enumerate(self.devices)}
        self.rank_of = {d: i for i, d in enumerate(self.devices)}
        self.streams: Dict[int, torch.cuda.Stream] = {d: torch.cuda.Stream(torch.device("cuda", d)) for d in self.devices}

    def send_recv(self, src: torch.Tensor, dst: torch.Tensor) -> torch.cuda.Event:
        """Copy ``src`` (device A) into ``dst`` (device B) on the side streams; the copy starts after the work
        currently queued on A's current stream.  Returns the event recorded on B's side stream."""
        lib = _load()
        a, b = src.device.index, dst.device.index
        nbytes = src.numel() * src.element_size()
        assert nbytes == dst.numel() * dst.element_size() and src.is_contiguous() and dst.is_contiguous()
        sa, sb = self.streams[a], self.streams[b]
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(src.device))
        sa.wait_event(ready)
        # dst must not be overwritten before earlier consumers on B are done with it
        free = torch.cuda.Event()
        free.record(torch.cuda.current_stream(dst.device))
        sb.wait_event(free)
        _check(lib.ncclGroupStart(), "ncclGroupStart")
        _check(lib.ncclSend(src.data_ptr(), nbytes, _NCCL_UINT8, self.rank_of[b], self.comms[a], sa.cuda_stream), "ncclSend")
        _check(lib.ncclRecv(dst.data_ptr(), nbytes, _NCCL_UINT8, self.rank_of[a], self.comms[b], sb.cuda_stream), "ncclRecv")
        _check(lib.ncclGroupEnd(), "ncclGroupEnd")
        src.record_stream(sa)
        dst.record_stream(sb)
        done = torch.cuda.Event()
        done.record(sb)
        return done

    def close(self) -> None:
        lib = _load()
        for c in self.comms.values():
            lib.ncclCommDestroy(c)
        self.comms.clear()
