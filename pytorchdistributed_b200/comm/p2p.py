"""Peer-to-peer communication runtime (one process per GPU, NVLink 5 / NVSwitch).

Symmetric memory - the same number of bytes on every rank, every copy addressable from every process - comes from
one of three providers, probed at start-up in this order (the one in use is reported by ``P2PCollectives.mode`` and
printed once by rank 0; ``PTDB200_P2P_MODE=nvls|vmm|ipc`` caps the choice):

``nvls``  ``cuMemCreate`` blocks exported as POSIX file descriptors, imported and mapped by every peer, PLUS one
          multicast object per block (``cuMulticastCreate`` / ``AddDevice`` / ``BindMem``, ``csrc/vmm.cu``): the
          all-reduce then runs ``multimem.ld_reduce`` / ``multimem.st`` - the reduction happens inside the NVSwitch;
``vmm``   the same blocks without the multicast object (driver or fabric manager refuses it): two-shot kernels;
``ipc``   ``cudaMalloc`` + CUDA IPC handles (the round-1 path), two-shot kernels.

File descriptors cross process boundaries over a unix datagram socket (``SCM_RIGHTS``); everything else of the
hand-shake (socket names, success flags) goes through the ``torch.distributed`` control plane.

``P2PCollectives`` - ``all_reduce_avg`` / ``broadcast`` / sharded collectives as single kernels of ``csrc/p2p.cu``
that synchronise through per-rank signal pads.  No NCCL call on any of these paths.

Serves what the reference gets from ``init_process_group("nccl")`` + DDP's internal all-reduce/broadcast
(``R/02 DDP基本概念/ddp_gpus.py:22,35,41``).
"""
from __future__ import annotations

import array
import bisect
import ctypes
import os
import socket
import struct
import uuid
import weakref
from typing import Dict, List, Optional, Tuple

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
_lib.register("ptd_vmm_caps", [])
_lib.register("ptd_vmm_create", [L, I, I, P, P])
_lib.register("ptd_vmm_import_peer", [I, I, I])
_lib.register("ptd_vmm_mc_create", [I, P])
_lib.register("ptd_vmm_mc_import", [I, I])
_lib.register("ptd_vmm_mc_add_device", [I])
_lib.register("ptd_vmm_mc_bind", [I])
_lib.register("ptd_vmm_mc_drop", [I])
_lib.register("ptd_vmm_ptrs", [I, P, P, P])
_lib.register("ptd_vmm_destroy", [I])
_lib.register("ptd_allreduce_mc", [P, P, P, I, I, L, I, F, I, P])
_lib.register("ptd_allreduce_wire_bf16", [P, P, P, P, P, I, I, L, F, I, P])
_lib.register("ptd_broadcast_mc", [P, P, P, P, I, I, I, L, I, P])
_lib.register("ptd_allreduce_fused_sgd", [P, P, P, P, I, I, L, F, I, P, P])

_LIVE: "weakref.WeakSet" = weakref.WeakSet()
MODES = ("nvls", "vmm", "ipc")


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


def _all_ok(ok: bool, group) -> bool:
    """True when every rank of the group reports success (collective)."""
    flags: List = [None] * dist.get_world_size(group)
    dist.all_gather_object(flags, bool(ok), group=group)
    return all(flags)


class FdChannel:
    """Passes open file descriptors between the ranks of a group (same host) over unix datagram sockets.

    Every rank binds one socket in the abstract namespace and learns the others' names through the process group.
    A message carries (tag, sender rank) plus exactly one descriptor; messages of a later exchange that overtake an
    earlier one are stashed by tag."""

    def __init__(self, group=None) -> None:
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.sock = socket.socket(socket.AF_UNIX, socket.SOCK_DGRAM)
        self.sock.bind("\0ptdb200-" + uuid.uuid4().hex)
        self.sock.settimeout(float(os.environ.get("PTDB200_FD_TIMEOUT_S", "60")))
        names: List = [None] * self.world
        dist.all_gather_object(names, self.sock.getsockname(), group=group)      # everyone is bound before anyone sends
        self.names = names
        self._stash: Dict[Tuple[int, int], int] = {}
        self._tag = 0

    def _send(self, dst: int, tag: int, fd: int) -> None:
        # (socket.send_fds of Python 3.12 drops its `address` argument, hence sendmsg itself)
        self.sock.sendmsg([struct.pack("ii", tag, self.rank)],
                          [(socket.SOL_SOCKET, socket.SCM_RIGHTS, array.array("i", [fd]))], 0, self.names[dst])

    def _recv(self, tag: int, src: int) -> int:
        while (tag, src) not in self._stash:
            msg, fds, _flags, _addr = socket.recv_fds(self.sock, 16, 4)
            t, s = struct.unpack("ii", msg[:8])
            if len(fds) != 1:
                raise RuntimeError(f"fd channel: message from rank {s} carried {len(fds)} descriptors")
            self._stash[(t, s)] = fds[0]
        return self._stash.pop((tag, src))

    def all_to_all(self, fd: int) -> List[Optional[int]]:
        """Give ``fd`` to every peer; returns the descriptor received from each rank (``None`` at our own index)."""
        self._tag += 1
        for r in range(self.world):
            if r != self.rank:
                self._send(r, self._tag, fd)
        return [None if r == self.rank else self._recv(self._tag, r) for r in range(self.world)]

    def from_root(self, fd: Optional[int], root: int = 0) -> Optional[int]:
        """``root`` gives ``fd`` to everyone else; returns the received descriptor (``None`` on the root)."""
        self._tag += 1
        if self.rank == root:
            for r in range(self.world):
                if r != root:
                    self._send(r, self._tag, fd)
            return None
        return self._recv(self._tag, root)

    def close(self) -> None:
        for fd in self._stash.values():
            os.close(fd)
        self._stash.clear()
        self.sock.close()


class _Region:
    """One symmetric allocation: ``ptrs[r]`` = this process's mapping of rank r's copy, ``mc`` = multicast mapping."""

    local_ptr: int
    nbytes: int
    ptrs: List[int]
    mc: int = 0

    def close(self) -> None:  # pragma: no cover - overridden
        raise NotImplementedError


class SymmetricBuffer(_Region):
    """``nbytes`` of ``cudaMalloc`` memory on every rank of ``group``, exported with CUDA IPC (fallback provider)."""

    def __init__(self, nbytes: int, device: torch.device, group=None) -> None:
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = device
        self.nbytes = (int(nbytes) + 255) // 256 * 256
        self.mc = 0
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

    def close(self, barrier: bool = True) -> None:
        """``barrier=False``: the caller has already synchronised the group (one barrier for many regions)."""
        if self._closed:
            return
        self._closed = True
        try:
            torch.cuda.synchronize(self.device)
            with torch.cuda.device(self.device):
                for p in self._opened:
                    _lib.call("ptd_p2p_ipc_close", p)
                if barrier and dist.is_initialized():
                    dist.barrier(group=self.group)      # nobody frees while a peer still has it mapped
                _lib.call("ptd_p2p_free", self.local_ptr)
        except Exception:  # noqa: BLE001 - teardown must not raise
            pass


class VmmBlock(_Region):
    """A symmetric block on the VMM API, optionally with its multicast (NVLS) mapping.  Collective constructor."""

    def __init__(self, nbytes: int, device: torch.device, group, chan: FdChannel, want_mc: bool) -> None:
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = device
        self.mc = 0
        self.block_id = 0
        self._closed = False
        err = None
        with torch.cuda.device(device):
            padded, fd = ctypes.c_longlong(0), ctypes.c_int(-1)
            bid = _lib.call_ret("ptd_vmm_create", int(nbytes), self.rank, self.world,
                                ctypes.cast(ctypes.pointer(padded), P), ctypes.cast(ctypes.pointer(fd), P))
            if bid <= 0:
                err = _lib.require().ptd_last_error().decode()
            if not _all_ok(bid > 0, group):
                if bid > 0:
                    os.close(fd.value)
                    _lib.call_ret("ptd_vmm_destroy", bid)
                raise RuntimeError(f"vmm_create failed on some rank ({err or 'not this one'})")
            self.block_id = bid
            self.nbytes = int(padded.value)
            theirs = chan.all_to_all(fd.value)
            os.close(fd.value)
            ok = True
            for r, pfd in enumerate(theirs):
                if pfd is None:
                    continue
                if ok and _lib.call_ret("ptd_vmm_import_peer", bid, r, pfd) != 0:
                    ok, err = False, _lib.require().ptd_last_error().decode()
                os.close(pfd)
            if not _all_ok(ok, group):
                _lib.call_ret("ptd_vmm_destroy", bid)
                raise RuntimeError(f"vmm_import_peer failed on some rank ({err or 'not this one'})")
            self.mc_error: Optional[str] = None
            if want_mc:
                self._setup_multicast(chan)
            uc = (ctypes.c_void_p * self.world)()
            mc = ctypes.c_void_p()
            _lib.call("ptd_vmm_ptrs", bid, ctypes.cast(uc, P), ctypes.cast(ctypes.pointer(mc), P), None)
            self.ptrs = [int(p) for p in uc]
            self.mc = int(mc.value or 0)
            self.local_ptr = self.ptrs[self.rank]
        self.bytes_tensor = tensor_from_ptr(self.local_ptr, self.nbytes, device)
        _LIVE.add(self)

    def _setup_multicast(self, chan: FdChannel) -> None:
        """All-or-nothing: any rank failing any step makes every rank drop the multicast object."""
        bid, lib = self.block_id, _lib.require()

        def step(ok: bool) -> bool:
            if not ok and self.mc_error is None:
                self.mc_error = lib.ptd_last_error().decode()
            return _all_ok(ok, self.group)

        fd = ctypes.c_int(-1)
        ok = True
        if self.rank == 0:
            ok = _lib.call_ret("ptd_vmm_mc_create", bid, ctypes.cast(ctypes.pointer(fd), P)) == 0
        if not step(ok):
            _lib.call_ret("ptd_vmm_mc_drop", bid)
            return
        got = chan.from_root(fd.value if self.rank == 0 else None, 0)
        if self.rank == 0:
            os.close(fd.value)
        else:
            ok = _lib.call_ret("ptd_vmm_mc_import", bid, got) == 0
            os.close(got)
        ok = ok and _lib.call_ret("ptd_vmm_mc_add_device", bid) == 0
        if not step(ok):                      # also the barrier the driver asks for: every device added before a bind
            _lib.call_ret("ptd_vmm_mc_drop", bid)
            return
        ok = _lib.call_ret("ptd_vmm_mc_bind", bid) == 0
        if not step(ok):
            _lib.call_ret("ptd_vmm_mc_drop", bid)

    def close(self, barrier: bool = True) -> None:
        if self._closed:
            return
        self._closed = True
        try:
            torch.cuda.synchronize(self.device)
            if barrier and dist.is_initialized():
                dist.barrier(group=self.group)          # no peer kernel still touches this block
            with torch.cuda.device(self.device):
                _lib.call_ret("ptd_vmm_destroy", self.block_id)
        except Exception:  # noqa: BLE001 - teardown must not raise
            pass


def probe_mode(group, device: torch.device) -> str:
    """Best provider every rank of the group supports (collective)."""
    cap = os.environ.get("PTDB200_P2P_MODE", "nvls")
    if cap not in MODES:
        raise RuntimeError(f"PTDB200_P2P_MODE={cap!r}: expected one of {MODES}")
    with torch.cuda.device(device):
        caps = _lib.call_ret("ptd_vmm_caps")
    allc: List = [None] * dist.get_world_size(group)
    dist.all_gather_object(allc, int(caps), group=group)
    both = allc[0]
    for c in allc:
        both &= c
    if cap == "nvls" and (both & 2):
        return "nvls"
    if cap in ("nvls", "vmm") and (both & 1):
        return "vmm"
    return "ipc"


class P2PCollectives:
    """Gradient averaging and broadcast over peer memory.  Same interface as ``ddp._TorchCollectives``.

    ``wire_dtype=torch.bfloat16``: fp32 tensors are all-reduced with bf16 on the wire (cast, reduction with fp32
    accumulation and cast back fused in one kernel) - half the NVLink bytes, results identical on every rank."""

    name = "p2p"

    def __init__(self, group, device: torch.device, wire_dtype: Optional[torch.dtype] = None,
                 arena_bytes: Optional[int] = None) -> None:
        if not _lib.available():
            raise RuntimeError(f"p2p backend needs the native library: {_lib.load_error()}")
        if wire_dtype not in (None, torch.float32, torch.bfloat16):
            raise RuntimeError(f"wire_dtype {wire_dtype}: only torch.bfloat16 (or None / float32 = off) is supported")
        self.group = group
        self.device = device
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        if self.world > 8:
            raise RuntimeError("p2p backend covers one NVSwitch domain (<= 8 ranks)")
        self.wire_dtype = torch.bfloat16 if wire_dtype == torch.bfloat16 else None
        self.blocks = int(os.environ.get("PTDB200_AR_BLOCKS", "0"))
        self.mc_barrier = os.environ.get("PTDB200_MC_BARRIER", "1") != "0"
        # Which all-reduce runs on multicast-capable memory.  Per direction of a port the switch path moves (1 + 1/N) x
        # the buffer, the two-shot path 2 (N-1)/N x: two ranks are better off peer-to-peer (1.0 vs 1.5), three break even,
        # four and more win through the switch (N = 8: 1.125 vs 1.75).  Measured at N = 2: 433 vs 684 us at 256 MB.
        self.algo = os.environ.get("PTDB200_AR_ALGO", "auto")
        if self.algo not in ("auto", "nvls", "two_shot"):
            raise RuntimeError(f"PTDB200_AR_ALGO={self.algo!r}: expected auto, nvls or two_shot")
        self.arena_bytes = int(arena_bytes or float(os.environ.get("PTDB200_ARENA_MB", "128")) * (1 << 20))
        self.stats = {"nvls": 0, "two_shot": 0, "wire_bf16": 0}
        self._regions: List[_Region] = []           # sorted by local_ptr
        self._bases: List[int] = []
        self._arena: Optional[Tuple[VmmBlock, int]] = None      # (block, bytes used)
        self._ptr_cache: Dict[Tuple[int, int], ctypes.Array] = {}
        self._wire_stage: Optional[torch.Tensor] = None
        self._chan: Optional[FdChannel] = None
        self.mode = probe_mode(group, device)
        self.mode_note = ""
        if self.mode != "ipc":
            self._chan = FdChannel(group)
            try:
                blk = VmmBlock(self.arena_bytes, device, group, self._chan, want_mc=self.mode == "nvls")
                if self.mode == "nvls" and not blk.mc:
                    self.mode, self.mode_note = "vmm", f"multicast refused: {blk.mc_error or 'by a peer'}"
                self._add_region(blk)
                self._arena = (blk, 0)
            except RuntimeError as e:            # collective failure: every rank takes this branch
                self.mode, self.mode_note = "ipc", f"VMM path refused: {e}"
        # Flags on one signal pad only grow (sequence counters live on the device next to the pad), so kernels
        # sharing a pad must be stream-ordered.  Two channels:
        # 0 = gradient all-reduce (DDP's communication stream), 1 = broadcasts (the compute stream).
        pad_bytes = _lib.call_ret("ptd_p2p_pad_bytes")
        self.pads = [self.alloc((pad_bytes + 255) // 256 * 256, torch.uint8, device) for _ in range(2)]
        if self.rank == 0 and os.environ.get("PTDB200_QUIET", "0") != "1":
            note = f" ({self.mode_note})" if self.mode_note else ""
            print(f"[ptdb200] p2p collectives: mode={self.mode}{note}, world={self.world}, "
                  f"wire={'bf16' if self.wire_dtype else 'native'}", flush=True)
        dist.barrier(group=group)

    # -- memory ---------------------------------------------------------------------------------------------------
    def _add_region(self, reg: _Region) -> None:
        i = bisect.bisect_left(self._bases, reg.local_ptr)
        self._bases.insert(i, reg.local_ptr)
        self._regions.insert(i, reg)

    def alloc(self, numel: int, dtype, device=None) -> torch.Tensor:
        """A zero-initialised symmetric tensor (collective: every rank calls with the same arguments)."""
        esz = torch.empty((), dtype=dtype).element_size()
        nbytes = max(numel * esz, 256)
        if self.mode == "ipc":
            sb = SymmetricBuffer(nbytes, self.device, self.group)
            self._add_region(sb)
            return sb.tensor(numel, dtype)
        need = (nbytes + 4095) // 4096 * 4096
        blk, used = self._arena
        if used + need > blk.nbytes:
            blk = VmmBlock(max(need, self.arena_bytes), self.device, self.group, self._chan, want_mc=self.mode == "nvls")
            if self.mode == "nvls" and not blk.mc:
                raise RuntimeError(f"multicast set-up failed for a later block: {blk.mc_error}")
            self._add_region(blk)
            used = 0
        self._arena = (blk, used + need)
        return blk.bytes_tensor[used:used + numel * esz].view(dtype)

    def _loc(self, t: torch.Tensor) -> Tuple[_Region, int]:
        ptr = t.data_ptr()
        i = bisect.bisect_right(self._bases, ptr) - 1
        if i >= 0:
            reg = self._regions[i]
            if ptr + t.numel() * t.element_size() <= reg.local_ptr + reg.nbytes:
                return reg, ptr - reg.local_ptr
        raise RuntimeError("p2p collective on a tensor that was not allocated by P2PCollectives.alloc()")

    def is_symmetric(self, t: torch.Tensor) -> bool:
        try:
            self._loc(t)
            return True
        except RuntimeError:
            return False

    def _ptrs(self, reg: _Region, off: int):
        key = (id(reg), off)
        arr = self._ptr_cache.get(key)
        if arr is None:
            arr = (ctypes.c_void_p * self.world)(*[p + off for p in reg.ptrs])
            self._ptr_cache[key] = arr
        return ctypes.cast(arr, P)

    def _pad(self, ch: int):
        """(unicast pad pointer array, multicast pad pointer or None) of channel ``ch``."""
        reg, off = self._loc(self.pads[ch])
        return self._ptrs(reg, off), (reg.mc + off if (reg.mc and self.mc_barrier) else None)

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    # -- all-reduce -----------------------------------------------------------------------------------------------
    def all_reduce_avg(self, flat: torch.Tensor, scale: Optional[float] = None, wire: Optional[bool] = None) -> None:
        """In place: ``flat = scale * sum over ranks`` (default ``scale = 1 / world``).  ``wire`` overrides whether
        an fp32 tensor travels as bf16 (default: the backend's ``wire_dtype``)."""
        reg, off = self._loc(flat)
        n = flat.numel()
        if flat.dtype not in (torch.float32, torch.bfloat16):
            raise RuntimeError(f"all_reduce_avg: unsupported dtype {flat.dtype}")
        is_bf16 = flat.dtype == torch.bfloat16
        scale = float(1.0 / self.world if scale is None else scale)
        pads, mc_pad = self._pad(0)
        use_wire = (self.wire_dtype is not None) if wire is None else bool(wire)
        if use_wire and not is_bf16 and self.world > 1:
            n8 = (n + 7) // 8 * 8                        # allocations are padded to 256 B, the tail is ours
            stage = self._stage(n8)
            sreg, soff = self._loc(stage)
            via_switch = sreg.mc and (self.algo == "nvls" or (self.algo == "auto" and self.world >= 3))
            _lib.call("ptd_allreduce_wire_bf16", flat, self._ptrs(sreg, soff), (sreg.mc + soff) if via_switch else None,
                      pads, mc_pad, self.rank, self.world, n8, scale, self.blocks, self._stream())
            self.stats["wire_bf16"] += 1
            return
        per = 8 if is_bf16 else 4
        n_pad = (n + per - 1) // per * per
        if reg.mc and self.world > 1 and (self.algo == "nvls" or (self.algo == "auto" and self.world >= 3)):
            _lib.call("ptd_allreduce_mc", reg.mc + off, pads, mc_pad, self.rank, self.world, n_pad, int(is_bf16), scale,
                      self.blocks, self._stream())
            self.stats["nvls"] += 1
            return
        _lib.call("ptd_allreduce_inplace", self._ptrs(reg, off), pads, self.rank, self.world, n_pad, int(is_bf16),
                  scale, self.blocks, self._stream())
        self.stats["two_shot"] += 1

    def bucket_comm(self, flat: torch.Tensor, sgd=None) -> "_lib.BucketComm":
        """The recipe of ``all_reduce_avg(flat)`` as a ``PtdBucketComm`` for the native reducer (fp32 buckets): the same
        algorithm choice, resolved once.  ``sgd``: an ``_lib.SgdEpilogue`` to fuse behind the reduction."""
        reg, off = self._loc(flat)
        n = flat.numel()
        c = _lib.BucketComm()
        c.rank, c.world, c.blocks = self.rank, self.world, self.blocks
        c.scale = 1.0 / self.world
        c.local = flat.data_ptr()
        via_switch = lambda r: bool(r.mc) and (self.algo == "nvls" or (self.algo == "auto" and self.world >= 3))  # noqa: E731
        for r in range(self.world):
            c.bufs[r] = reg.ptrs[r] + off
        preg, poff = self._loc(self.pads[0])
        for r in range(self.world):
            c.pads[r] = preg.ptrs[r] + poff
        c.mc_pad = (preg.mc + poff) if (preg.mc and self.mc_barrier) else None
        if self.wire_dtype is not None and sgd is None:
            n8 = (n + 7) // 8 * 8
            stage = self._stage(n8)
            sreg, soff = self._loc(stage)
            for r in range(self.world):
                c.stages[r] = sreg.ptrs[r] + soff
            c.mc_stage = (sreg.mc + soff) if via_switch(sreg) else None
            c.algo, c.numel = 3, n8
        else:
            c.numel = (n + 3) // 4 * 4
            if via_switch(reg):
                c.algo, c.mc_buf = 2, reg.mc + off
            else:
                c.algo = 1
        if sgd is not None:
            c.fuse_sgd, c.sgd = 1, sgd
        return c

    def count(self, c: "_lib.BucketComm") -> None:
        self.stats[{1: "two_shot", 2: "nvls", 3: "wire_bf16"}[c.algo]] += 1

    def _stage(self, numel: int) -> torch.Tensor:
        """bf16 staging buffer of the wire format, grown collectively (every rank sees the same request sizes)."""
        if self._wire_stage is None or self._wire_stage.numel() < numel:
            self._wire_stage = self.alloc(numel, torch.bfloat16, self.device)
        return self._wire_stage[:numel]

    def reserve_wire(self, numel: int) -> None:
        """Allocate the bf16 staging buffer up front (allocation is collective and synchronises: not inside backward)."""
        if self.wire_dtype is not None:
            self._stage((numel + 7) // 8 * 8)

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
        reg, off = self._loc(flat)
        if op in ("reduce_scatter", "reduce", "all_reduce_oneshot") and flat.dtype != torch.float32:
            raise RuntimeError(f"{op}: fp32 only (got {flat.dtype})")
        nbytes = (flat.numel() * flat.element_size() + 15) // 16 * 16      # allocations are padded to 256 B
        _lib.call("ptd_shard_collective", self._ptrs(reg, off), self._pad(1)[0], self.rank, self.world, root,
                  _OPS[op], nbytes, float(scale), self.blocks, self._stream())

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
        if not self.is_symmetric(flat):
            # not symmetric memory (e.g. a temporary during construction): stage through a symmetric buffer that is
            # kept and reused for the next such call
            stage = getattr(self, "_bcast_stage", None)
            if stage is None or stage.numel() < flat.numel() * flat.element_size():
                stage = self._bcast_stage = self.alloc(flat.numel() * flat.element_size(), torch.uint8, self.device)
            view = stage[:flat.numel() * flat.element_size()].view(flat.dtype)
            view.copy_(flat.reshape(-1))
            self.broadcast(view, src)
            flat.copy_(view.view(flat.shape))
            return
        reg, off = self._loc(flat)
        nbytes = (flat.numel() * flat.element_size() + 15) // 16 * 16
        pads, mc_pad = self._pad(1)
        if reg.mc and self.world > 1:
            _lib.call("ptd_broadcast_mc", self._ptrs(reg, off), reg.mc + off, pads, mc_pad, self.rank, self.world, src,
                      nbytes, self.blocks, self._stream())
            return
        _lib.call("ptd_broadcast_inplace", self._ptrs(reg, off), pads, self.rank, self.world, src, nbytes, self.blocks,
                  self._stream())

    def barrier(self) -> None:
        _lib.call("ptd_p2p_barrier", self._pad(1)[0], self.rank, self.world, self._stream())

    def close(self) -> None:
        """Collective: every rank calls it.  ONE group barrier, then every region is released locally."""
        try:
            torch.cuda.synchronize(self.device)
            if dist.is_initialized() and self._regions:
                dist.barrier(group=self.group)
        except Exception:  # noqa: BLE001 - teardown must not raise
            pass
        for reg in self._regions:
            reg.close(barrier=False)
        self._regions.clear()
        self._bases.clear()
        self._ptr_cache.clear()
        self._arena = None
        self._wire_stage = None
        if self._chan is not None:
            self._chan.close()
            self._chan = None


def shutdown() -> None:
    """Release every live symmetric allocation (called from ``destroy_process_group``).

    Exactly ONE barrier, whatever this rank still holds: which regions are alive here depends on when Python's cyclic
    garbage collector ran, and that differs between ranks - a barrier per region would leave the rank that holds more of
    them waiting for a peer that has already left."""
    regs = [r for r in list(_LIVE) if not getattr(r, "_closed", True)]
    try:
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        if dist.is_initialized() and dist.get_world_size() > 1 and dist.get_backend() != "gloo":
            dist.barrier()
    except Exception:  # noqa: BLE001 - teardown must not raise
        pass
    for reg in regs:
        reg.close(barrier=False)
