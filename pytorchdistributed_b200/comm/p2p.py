"""Peer-to-peer communication runtime (one process per GPU, NVLink/NVSwitch).

* ``SymmetricBuffer`` - the same-sized ``cudaMalloc`` block on every rank, exported with CUDA IPC and
  mapped by every peer (handles travel through the ``torch.distributed`` control plane once, at
  allocation time).  The local block is exposed as a torch tensor; gradient buckets and the flat
  BatchNorm-buffer block of ``DistributedDataParallel`` live in such buffers.
* ``P2PCollectives``  - ``all_reduce_avg`` / ``broadcast`` as single kernels of ``csrc/p2p.cu`` that read
  and write the peers' blocks directly and synchronise through per-rank signal pads.  No NCCL call.

Serves what the reference gets from ``init_process_group("nccl")`` + DDP's internal all-reduce/broadcast
(``R/02 DDP基本概念/ddp_gpus.py:22,35,41``).
"""
from __future__ import annotations

import ctypes
import os
import weakref
from typing import List, Optional

import torch
import torch.distributed as dist

from .. import _lib

P, I, L, F, U = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_uint
_lib.register("ptd_p2p_malloc", [L, P])
_lib.register("ptd_p2p_free", [P])
_lib.register("ptd_p2p_ipc_handle_size", [])
_lib.register("ptd_p2p_ipc_get", [P, P])
_lib.register("ptd_p2p_ipc_open", [P, P])
_lib.register("ptd_p2p_ipc_close", [P])
_lib.register("ptd_p2p_enable_peer_access", [I, I])
_lib.register("ptd_p2p_pad_bytes", [])
_lib.register("ptd_p2p_max_blocks", [])
_lib.register("ptd_allreduce_inplace", [P, P, I, I, L, I, F, I, P])
_lib.register("ptd_broadcast_inplace", [P, P, I, I, I, L, I, P])
_lib.register("ptd_p2p_barrier", [P, I, I, P])
_lib.register("ptd_shard_collective", [P, P, I, I, I, I, L, F, I, P])
_OPS = {"all_gather": 0, "gather": 1, "scatter": 2, "reduce_scatter": 3, "reduce": 4, "all_reduce_oneshot": 5}
_lib.register("ptd_peer_copy", [P, P, L, P])
_lib.register("ptd_reduce_sum_f32", [P, P, I, L, F, I, P])
_lib.register("ptd_peer_copy_bulk", [P, P, L, I, P])

_LIVE: "weakref.WeakSet[SymmetricBuffer]" = weakref.WeakSet()


class _RawCuda:
    """Minimal ``__cuda_array_interface__`` holder so a raw device pointer becomes a torch tensor."""

    def __init__(self, ptr: int, nbytes: int) -> None:
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


def raw_malloc(nbytes: int) -> int:
    out = ctypes.c_void_p()
    _lib.call("ptd_p2p_malloc", nbytes, ctypes.cast(ctypes.pointer(out), ctypes.c_void_p))
    return int(out.value)


def tensor_from_ptr(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    with torch.cuda.device(device):
        return torch.as_tensor(_RawCuda(ptr, nbytes), device=device)


class SymmetricBuffer:
    """``nbytes`` of device memory on every rank of ``group``, each mapped into every process."""

    def __init__(self, nbytes: int, device: torch.device, group=None) -> None:
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = device
        self.nbytes = (int(nbytes) + 255) // 256 * 256
        with torch.cuda.device(device):
            self.local_ptr = raw_malloc(self.nbytes)
            hs = _lib.call_ret("ptd_p2p_ipc_handle_size")
            hbuf = (ctypes.c_ubyte * hs)()
            _lib.call("ptd_p2p_ipc_get", self.local_ptr, ctypes.cast(hbuf, ctypes.c_void_p))
            handles: List = [None] * self.world
            dist.all_gather_object(handles, bytes(hbuf), group=group)
            self.ptrs: List[int] = []
            self._opened: List[int] = []
            for r, h in enumerate(handles):
                if r == self.rank:
                    self.ptrs.append(self.local_ptr)
                    continue
                out = ctypes.c_void_p()
                hb = (ctypes.c_ubyte * hs).from_buffer_copy(h)
                _lib.call("ptd_p2p_ipc_open", ctypes.cast(hb, ctypes.c_void_p),
                          ctypes.cast(ctypes.pointer(out), ctypes.c_void_p))
                self.ptrs.append(int(out.value))
                self._opened.append(int(out.value))
        self.ptr_array = (ctypes.c_void_p * self.world)(*self.ptrs)      # host array handed to the entry points
        self.bytes_tensor = tensor_from_ptr(self.local_ptr, self.nbytes, device)
        self._closed = False
        _LIVE.add(self)

    def tensor(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        need = numel * torch.empty((), dtype=dtype).element_size()
        assert need <= self.nbytes
        return self.bytes_tensor[:need].view(dtype)

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        try:
            torch.cuda.synchronize(self.device)
            with torch.cuda.device(self.device):
                for p in self._opened:
                    _lib.call("ptd_p2p_ipc_close", p)
                if dist.is_initialized():
                    dist.barrier(group=self.group)      # nobody frees while a peer still has it mapped
                _lib.call("ptd_p2p_free", self.local_ptr)
        except Exception:  # noqa: BLE001 - teardown must not raise
            pass


class P2PCollectives:
    """Gradient averaging and broadcast over peer memory.  Same interface as ``ddp._TorchCollectives``."""

    name = "p2p"

    def __init__(self, group, device: torch.device, wire_dtype: Optional[torch.dtype] = None) -> None:
        if not _lib.available():
            raise RuntimeError(f"p2p backend needs the native library: {_lib.load_error()}")
        self.group = group
        self.device = device
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        if self.world > 8:
            raise RuntimeError("p2p backend covers one NVSwitch domain (<= 8 ranks)")
        # Flags on one signal pad only grow (sequence counters live on the device next to the pad), so kernels
        # sharing a pad must be stream-ordered.  Two channels:
        # 0 = gradient all-reduce (DDP's communication stream), 1 = broadcasts (the compute stream).
        self.pads = [SymmetricBuffer(_lib.call_ret("ptd_p2p_pad_bytes"), device, group) for _ in range(2)]
        self.blocks = int(os.environ.get("PTDB200_AR_BLOCKS", "0"))
        self._bufs: List[SymmetricBuffer] = []
        self._by_ptr = {}
        dist.barrier(group=group)

    def alloc(self, numel: int, dtype, device) -> torch.Tensor:
        esz = torch.empty((), dtype=dtype).element_size()
        sb = SymmetricBuffer(max(numel * esz, 256), self.device, self.group)
        t = sb.tensor(numel, dtype)
        self._bufs.append(sb)
        self._by_ptr[t.data_ptr()] = sb
        return t

    def _sym(self, flat: torch.Tensor) -> SymmetricBuffer:
        sb = self._by_ptr.get(flat.data_ptr())
        if sb is None:
            raise RuntimeError("p2p collective on a tensor that was not allocated by P2PCollectives.alloc()")
        return sb

    def all_reduce_avg(self, flat: torch.Tensor, scale: Optional[float] = None) -> None:
        sb = self._sym(flat)
        n = flat.numel()
        is_bf16 = flat.dtype == torch.bfloat16
        if flat.dtype not in (torch.float32, torch.bfloat16):
            raise RuntimeError(f"all_reduce_avg: unsupported dtype {flat.dtype}")
        per = 8 if is_bf16 else 4
        n_pad = (n + per - 1) // per * per               # buffers are padded to 256 B, the tail is ours
        _lib.call("ptd_allreduce_inplace", ctypes.cast(sb.ptr_array, P), ctypes.cast(self.pads[0].ptr_array, P),
                  self.rank, self.world, n_pad, int(is_bf16), float(1.0 / self.world if scale is None else scale),
                  self.blocks, torch.cuda.current_stream(self.device).cuda_stream)

    # -- sharded collectives (in place on a tensor from alloc(); shard s = rank s's part) ----------------------
    def shard(self, flat: torch.Tensor, rank: Optional[int] = None) -> torch.Tensor:
        """The slice of ``flat`` that belongs to ``rank`` (default: this rank) in the sharded collectives:
        ``ceil(nbytes / 16 / world)`` 16-byte vectors per rank, the last shard may be short or empty."""
        rank = self.rank if rank is None else rank
        esz = flat.element_size()
        n16 = (flat.numel() * esz + 15) // 16
        per = (n16 + self.world - 1) // self.world * 16 // esz
        return flat[min(flat.numel(), per * rank):min(flat.numel(), per * (rank + 1))]

    def _sharded(self, op: str, flat: torch.Tensor, root: int = 0, scale: float = 1.0) -> None:
        sb = self._sym(flat)
        if op in ("reduce_scatter", "reduce", "all_reduce_oneshot") and flat.dtype != torch.float32:
            raise RuntimeError(f"{op}: fp32 only (got {flat.dtype})")
        nbytes = (flat.numel() * flat.element_size() + 15) // 16 * 16      # buffers are padded to 256 B
        _lib.call("ptd_shard_collective", ctypes.cast(sb.ptr_array, P), ctypes.cast(self.pads[1].ptr_array, P),
                  self.rank, self.world, root, _OPS[op], nbytes, float(scale), self.blocks,
                  torch.cuda.current_stream(self.device).cuda_stream)

    def all_gather(self, flat: torch.Tensor) -> None:
        """Every rank wrote ``shard(flat)``; afterwards every rank holds all shards (dist.all_gather_into_tensor)."""
        self._sharded("all_gather", flat)

    def gather(self, flat: torch.Tensor, dst: int = 0) -> None:
        """Like :meth:`all_gather`, but only ``dst`` ends up with all shards (dist.gather)."""
        self._sharded("gather", flat, dst)

    def scatter(self, flat: torch.Tensor, src: int = 0) -> None:
        """Rank r receives ``shard(flat, r)`` of ``src``'s tensor into the same slice of its own (dist.scatter)."""
        self._sharded("scatter", flat, src)

    def reduce_scatter(self, flat: torch.Tensor, scale: float = 1.0) -> None:
        """``shard(flat)`` becomes ``scale *`` the sum over ranks of that slice (dist.reduce_scatter_tensor);
        the rest of ``flat`` keeps this rank's own contribution."""
        self._sharded("reduce_scatter", flat, 0, scale)

    def reduce(self, flat: torch.Tensor, dst: int = 0, scale: float = 1.0) -> None:
        """``dst``'s tensor becomes ``scale *`` the sum over ranks (dist.reduce); the others keep theirs."""
        self._sharded("reduce", flat, dst, scale)

    def all_reduce_oneshot(self, flat: torch.Tensor, scale: Optional[float] = None) -> None:
        """All-reduce for small fp32 messages (<= 4 MiB): one NVLink round trip instead of two."""
        self._sharded("all_reduce_oneshot", flat, 0, 1.0 / self.world if scale is None else scale)

    def broadcast(self, flat: torch.Tensor, src: int = 0) -> None:
        sb = self._by_ptr.get(flat.data_ptr())
        if sb is None:
            # not symmetric memory (e.g. a temporary during construction): stage through a symmetric buffer
            stage = self.alloc(flat.numel(), flat.dtype, self.device)
            stage.copy_(flat)
            self.broadcast(stage, src)
            flat.copy_(stage)
            return
        nbytes = (flat.numel() * flat.element_size() + 15) // 16 * 16
        _lib.call("ptd_broadcast_inplace", ctypes.cast(sb.ptr_array, P), ctypes.cast(self.pads[1].ptr_array, P),
                  self.rank, self.world, src, nbytes, self.blocks,
                  torch.cuda.current_stream(self.device).cuda_stream)

    def barrier(self) -> None:
        _lib.call("ptd_p2p_barrier", ctypes.cast(self.pads[1].ptr_array, P), self.rank, self.world,
                  torch.cuda.current_stream(self.device).cuda_stream)

    def close(self) -> None:
        for sb in self._bufs:
            sb.close()
        for pad in self.pads:
            pad.close()
        self._bufs.clear()
        self._by_ptr.clear()


def shutdown() -> None:
    """Release every live symmetric buffer (called from ``destroy_process_group``)."""
    for sb in list(_LIVE):
        sb.close()
