"""``DistributedDataParallel`` - one process per GPU, bucketed gradient averaging overlapped with backward.

Surface and semantics follow what the reference uses (``DDP(model, device_ids=[gpu_id])`` at
``R/02 DDP基本概念/ddp_gpus.py:35`` / ``ddp_gpus_torchrun.py:33``; inner module under ``.module`` so
``state_dict()`` keys carry the ``module.`` prefix):

* construction: parameter shapes are checked across ranks, then rank 0's parameters and buffers win;
* gradients live as views into flat per-bucket buffers (first bucket 1 MiB, then 25 MiB, reverse
  registration order; rebuilt once in observed ready order) - flatten is free;
* a post-accumulate-grad hook per parameter tells the native reducer (``csrc/reducer.cpp``) the
  gradient is ready; a bucket whose last gradient arrived is averaged across ranks right away on a
  communication stream, overlapping the rest of backward;
* ``no_sync()`` skips the reduction (local accumulation); buffers (BatchNorm statistics) are
  re-broadcast from rank 0, as one flat message, before every forward that follows a gradient-synchronised
  forward (torch's ``require_forward_param_sync`` rule: every training step, and the first evaluation pass
  after training).

What changes underneath: on CUDA with >1 rank the averaging is our own kernel over NVLink (``comm/p2p.py``):
through the NVSwitch (``multimem.ld_reduce`` / ``multimem.st``) where the box exposes multicast objects, a two-shot
peer-to-peer reduction otherwise - the 1/world scale fused in, and with ``wire_dtype=torch.bfloat16`` the fp32
buckets travel as bf16 (cast, reduction and cast back in the same kernel).  No NCCL call on that path;
``comm="nccl"``/``"gloo"`` selects the ``torch.distributed`` collective instead (baseline, and the CPU plumbing path).

Gradients reach the buckets without an extra pass where the producing kernel allows it: the module registers as the
*gradient sink* of ``ops.functional`` and the residual blocks' weight-gradient kernels reduce straight into the bucket
views (``_Bottleneck.backward``), then report their parameters ready themselves (``mark_ready``).
"""
from __future__ import annotations

import contextlib
import ctypes
import os
import time
from typing import Dict, List, Optional

import torch
import torch.distributed as dist
from torch import nn

from .. import _lib
from ..ops import functional as _F
from ..utils.nvtx import nvtx_range

_MiB = 1024 * 1024

_lib.register("ptd_plan_buckets", [ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_void_p])
_lib.register("ptd_reducer_create", [ctypes.c_void_p, ctypes.c_int, ctypes.c_int])
_lib.register("ptd_reducer_bind_bucket", [ctypes.c_int, ctypes.c_int, ctypes.c_void_p])
_lib.register("ptd_reducer_mark_ready_launch", [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p])
_lib.register("ptd_reducer_finalize", [ctypes.c_int, ctypes.c_void_p])
_lib.register("ptd_reducer_mark_ready", [ctypes.c_int, ctypes.c_int])
_lib.register("ptd_reducer_all_done", [ctypes.c_int])
_lib.register("ptd_reducer_ready_order", [ctypes.c_int, ctypes.c_void_p, ctypes.c_int])
_lib.register("ptd_reducer_incomplete", [ctypes.c_int, ctypes.c_void_p, ctypes.c_int])
_lib.register("ptd_reducer_reset", [ctypes.c_int])
_lib.register("ptd_reducer_destroy", [ctypes.c_int])


def plan_buckets(sizes_bytes: List[int], first_cap: int, cap: int) -> List[int]:
    """Bucket index per tensor, tensors taken in the given order (native planner)."""
    n = len(sizes_bytes)
    if n == 0:
        return []
    sizes = (ctypes.c_longlong * n)(*sizes_bytes)
    out = (ctypes.c_int * n)()
    nb = _lib.call_ret("ptd_plan_buckets", ctypes.cast(sizes, ctypes.c_void_p), n, first_cap, cap,
                       ctypes.cast(out, ctypes.c_void_p))
    if nb < 0:
        raise RuntimeError("ptd_plan_buckets failed")
    return list(out)


class _TorchCollectives:
    """Gradient averaging / broadcast through torch.distributed (gloo on CPU, NCCL baseline on GPU)."""

    name = "torch"

    def __init__(self, group) -> None:
        self.group = group
        self.world = dist.get_world_size(group)

    def alloc(self, numel: int, dtype, device) -> torch.Tensor:
        return torch.zeros(numel, dtype=dtype, device=device)

    def all_reduce_avg(self, flat: torch.Tensor) -> None:
        if flat.is_cuda:
            dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group)
        else:
            dist.all_reduce(flat, group=self.group)
            flat.div_(self.world)

    def broadcast(self, flat: torch.Tensor, src: int = 0) -> None:
        dist.broadcast(flat, src=dist.get_global_rank(self.group, src) if self.group is not None else src,
                       group=self.group)

    def close(self) -> None:
        pass


class _Bucket:
    __slots__ = ("flat", "params", "views", "offsets", "ready_event", "done_event", "launched", "t0", "t1")

    def __init__(self) -> None:
        self.params: List[nn.Parameter] = []
        self.views: List[torch.Tensor] = []
        self.offsets: List[int] = []
        self.flat: Optional[torch.Tensor] = None
        self.ready_event = None
        self.done_event = None
        self.launched = False
        self.t0 = self.t1 = None            # communication timing (enable_comm_timing): events on CUDA, seconds on CPU


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, dim: int = 0,
                 broadcast_buffers: bool = True, process_group=None, bucket_cap_mb: float = 25,
                 first_bucket_mb: float = 1, find_unused_parameters: bool = False,
                 gradient_as_bucket_view: bool = True, comm: str = "auto", wire_dtype: Optional[torch.dtype] = None,
                 rebuild_buckets: bool = True, direct_grads: bool = True) -> None:
        super().__init__()
        if not dist.is_initialized():
            raise RuntimeError("DistributedDataParallel needs an initialised process group (comm.ddp_setup)")
        # find_unused_parameters: parameters that received no gradient in a backward are reported ready at its end
        # (their bucket views hold zeros), so their buckets are still averaged - every rank must leave the SAME
        # parameters unused (a static sub-graph), the rank-divergent case of torch's bitmap exchange is not covered.
        self.find_unused_parameters = bool(find_unused_parameters)
        self._comm_hook = None
        self._time_comm = False
        self._comm_ms: List[float] = []
        self._timed_order: List[_Bucket] = []
        self.module = module
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group)
        self.rank = dist.get_rank(process_group)
        # PTDB200_DDP_BCAST=0 / PTDB200_DDP_SKIP_ALLREDUCE=1: control runs that attribute step time to the per-forward
        # buffer broadcast / the gradient all-reduce (results are then wrong on purpose: replicas drift apart)
        self.broadcast_buffers = broadcast_buffers and os.environ.get("PTDB200_DDP_BCAST", "1") != "0"
        self._skip_allreduce = os.environ.get("PTDB200_DDP_SKIP_ALLREDUCE", "0") == "1"
        self.require_backward_grad_sync = True
        self._require_forward_param_sync = True
        self.bucket_cap = int(bucket_cap_mb * _MiB)
        self.first_cap = int(first_bucket_mb * _MiB)
        self._want_rebuild = rebuild_buckets
        self._rebuilt = False
        self._params: List[nn.Parameter] = [p for p in module.parameters() if p.requires_grad]
        if not self._params:
            raise RuntimeError("DistributedDataParallel: module has no parameter that requires grad")
        self.device = self._params[0].device
        if device_ids is not None and self.device.type == "cuda":
            want = torch.device("cuda", device_ids[0]) if isinstance(device_ids[0], int) else torch.device(device_ids[0])
            if want.index is not None and want != self.device:
                raise RuntimeError(f"module is on {self.device} but device_ids={device_ids}")
        self.backend = self._make_backend(comm, wire_dtype)
        self._comm_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._verify_shapes()
        self._flatten_buffers()
        self._sync_module_states()
        self._index_of: Dict[int, int] = {id(p): i for i, p in enumerate(self._params)}
        self._reducer_id = 0
        self._buckets: List[_Bucket] = []
        self._native = False
        self._fused_opt = None
        self._grad_region: Optional[torch.Tensor] = None      # one symmetric allocation all buckets are carved from
        self._build_buckets(list(reversed(range(len(self._params)))))
        self._callback_queued = False
        self._hooks = [p.register_post_accumulate_grad_hook(self._grad_ready) for p in self._params]
        self.stats = {"allreduce_launches": 0, "bytes_reduced": 0, "iterations": 0, "direct_grads": 0}
        self.direct_grads = direct_grads and os.environ.get("PTDB200_DDP_DIRECT", "1") != "0"
        if self.direct_grads and self.device.type == "cuda":
            _F.set_grad_sink(self)            # the most recently built wrapper receives gradients directly

    # ------------------------------------------------------------------ construction helpers
    def _make_backend(self, comm: str, wire_dtype):
        comm = os.environ.get("PTDB200_DDP_COMM", comm)
        if comm == "auto":
            comm = "p2p" if (self.device.type == "cuda" and self.world_size > 1) else "torch"
        if comm == "p2p":
            if wire_dtype is None and os.environ.get("PTDB200_DDP_WIRE", "") == "bf16":
                wire_dtype = torch.bfloat16
            from ..comm import p2p
            return p2p.P2PCollectives(self.process_group, self.device, wire_dtype=wire_dtype)
        return _TorchCollectives(self.process_group)

    def _verify_shapes(self) -> None:
        mine = [tuple(p.shape) for p in self._params]
        allv: List = [None] * self.world_size
        dist.all_gather_object(allv, mine, group=self.process_group)
        for r, other in enumerate(allv):
            if other != mine:
                raise RuntimeError(f"parameter shapes differ between rank {self.rank} and rank {r}")

    def _flatten_buffers(self) -> None:
        """Re-point every floating buffer at a slice of one flat tensor (same for integer buffers) so the
        per-forward buffer broadcast is one message instead of one per BatchNorm tensor."""
        self._buf_flats: List[torch.Tensor] = []
        groups: Dict[torch.dtype, List] = {}
        for mod in self.module.modules():
            for name, buf in mod._buffers.items():
                if buf is not None and name not in mod._non_persistent_buffers_set:
                    groups.setdefault(buf.dtype, []).append((mod, name, buf))
        for dtype, items in groups.items():
            total = sum(b.numel() for _, _, b in items)
            flat = self.backend.alloc(total, dtype, items[0][2].device)
            off = 0
            for mod, name, buf in items:
                n = buf.numel()
                view = flat[off:off + n].view(buf.shape)
                view.copy_(buf)
                mod._buffers[name] = view
                off += n
            self._buf_flats.append(flat)

    def _sync_module_states(self) -> None:
        """Rank 0's parameters and buffers win (torch DDP does the same at construction)."""
        if self.world_size == 1:
            return
        with torch.no_grad():
            chunks = [self._storage_view(p) for p in self.module.parameters()]
            by_dtype: Dict[torch.dtype, List[torch.Tensor]] = {}
            for c in chunks:
                by_dtype.setdefault(c.dtype, []).append(c)
            for _, group in by_dtype.items():
                flat = torch.cat(group)
                self.backend.broadcast(flat, 0)
                off = 0
                for c in group:
                    c.copy_(flat[off:off + c.numel()])
                    off += c.numel()
            for bf in self._buf_flats:
                self.backend.broadcast(bf, 0)

    @staticmethod
    def _storage_view(p: torch.Tensor) -> torch.Tensor:
        """1-D view of a dense tensor in memory order (so channels_last parameters need no permute)."""
        t = p.detach()
        return torch.as_strided(t, (t.numel(),), (1,), t.storage_offset())

    def _build_buckets(self, order: List[int]) -> None:
        sizes = [self._params[i].numel() * self._params[i].element_size() for i in order]
        assign = plan_buckets(sizes, self.first_cap, self.bucket_cap)
        n_buckets = max(assign) + 1
        # a rebuild re-carves the region the old views live in: keep what they hold (accumulated gradients) aside
        old_grads = {id(p): (p.grad.clone() if (p.grad is not None and self._grad_region is not None) else p.grad)
                     for p in self._params}
        buckets = [_Bucket() for _ in range(n_buckets)]
        bucket_of = [0] * len(self._params)
        for pos, pi in enumerate(order):
            buckets[assign[pos]].params.append(self._params[pi])
            bucket_of[pi] = assign[pos]
        totals = []
        for b in buckets:
            total, offs = 0, []
            for p in b.params:
                total = (total + 3) // 4 * 4          # 16-byte alignment of every fp32 view
                offs.append(total)
                total += p.numel()
            total = (total + 63) // 64 * 64           # buckets start 256-byte aligned inside the region
            b.offsets = offs
            totals.append(total)
        dtype = self._params[0].dtype
        if any(p.dtype != dtype for p in self._params):
            raise RuntimeError("DistributedDataParallel: parameters must share one dtype")
        need = sum(totals)
        if self._grad_region is None or self._grad_region.numel() < need:
            # slack: a rebuild in ready order changes the per-bucket padding a little and must fit the same region
            # (a second region would double the symmetric memory held until shutdown)
            self._grad_region = self.backend.alloc(need + 64 * (len(self._params) + 64), dtype, self.device)
        else:
            self._grad_region.zero_()
        start = 0
        for b, total in zip(buckets, totals):
            b.flat = self._grad_region[start:start + total]
            start += total
            for p, off in zip(b.params, b.offsets):
                view = torch.as_strided(b.flat, p.shape, p.stride(), b.flat.storage_offset() + off)
                g = old_grads[id(p)]
                if g is not None:
                    view.copy_(g)
                p.grad = view
                b.views.append(view)
            if self.device.type == "cuda":
                b.ready_event = torch.cuda.Event()
                b.done_event = torch.cuda.Event()
        if hasattr(self.backend, "reserve_wire"):
            self.backend.reserve_wire(max(totals))
        self._buckets = buckets
        self._bucket_of = bucket_of
        arr = (ctypes.c_int * len(bucket_of))(*bucket_of)
        if self._reducer_id:
            _lib.call_ret("ptd_reducer_destroy", self._reducer_id)
        self._reducer_id = _lib.call_ret("ptd_reducer_create", ctypes.cast(arr, ctypes.c_void_p), len(bucket_of),
                                         n_buckets)
        if self._reducer_id <= 0:
            raise RuntimeError("ptd_reducer_create failed")
        self._view_of = {id(p): v for b in buckets for p, v in zip(b.params, b.views)}
        self._bucket_starts = [b.flat.storage_offset() - self._grad_region.storage_offset() for b in buckets]
        if getattr(self, "_fused_opt", None) is not None:
            self._rehome_state()               # the flat optimizer state follows the (new) bucket layout
        self._bind_native()

    def _bind_native(self) -> None:
        """Hand every bucket's reduction recipe to the native reducer: from then on marking a parameter ready is one
        native call that also issues the finished bucket's all-reduce (events, stream switch, kernel)."""
        self._native = False
        self._bucket_comms = []
        if (self.world_size == 1 or self._skip_allreduce or self._comm_stream is None or self._comm_hook is not None
                or self._time_comm or not hasattr(self.backend, "bucket_comm")
                or os.environ.get("PTDB200_DDP_NATIVE", "1") == "0"):
            return
        fused = getattr(self, "_fused_opt", None) is not None
        for bi, b in enumerate(self._buckets):
            sgd = self._sgd_epilogue(bi) if fused else None
            c = self.backend.bucket_comm(b.flat, sgd=sgd)
            _lib.call("ptd_reducer_bind_bucket", self._reducer_id, bi, ctypes.addressof(c))
            self._bucket_comms.append(c)
        self._native = True

    # ------------------------------------------------------------------ optimizer step fused behind the reduction
    def fuse_optimizer(self, optimizer) -> bool:
        """Let the all-reduce kernel of every bucket apply ``optimizer``'s SGD step to that bucket's parameters as its
        epilogue (``csrc/p2p.cu: sgd_epilogue``): the communication kernel becomes the last touch of each gradient, and
        ``optimizer.step()`` turns into bookkeeping.  Parameters, momentum buffers and the bf16 weight shadows are
        re-homed into flat arrays that mirror the bucket layout.  Needs our fused ``SGD`` (one parameter group), the
        peer-to-peer backend with fp32 on the wire and more than one rank; returns ``False`` (and changes nothing)
        otherwise.  Between ``backward()`` and ``step()`` the parameters are then already updated - gradient clipping or
        a skipped ``step()`` are not possible in this mode."""
        from ..optim import SGD
        ok = (isinstance(optimizer, SGD) and len(optimizer.param_groups) == 1 and self.world_size > 1
              and hasattr(self.backend, "bucket_comm") and getattr(self.backend, "wire_dtype", None) is None
              and self.device.type == "cuda" and not self._skip_allreduce
              and {id(p) for p in optimizer.param_groups[0]["params"]} == {id(p) for p in self._params})
        if not ok:
            return False
        self._fused_opt = optimizer
        self._fused_stepped = False
        self._rehome_state()
        self._bind_native()
        optimizer._ddp_fused = self
        return True

    def _rehome_state(self) -> None:
        opt = self._fused_opt
        group = opt.param_groups[0]
        n = self._grad_region.numel()
        with torch.no_grad():
            new_p = torch.zeros(n, dtype=torch.float32, device=self.device)
            new_s = torch.zeros(n, dtype=torch.bfloat16, device=self.device)
            new_m = torch.zeros(n, dtype=torch.float32, device=self.device) if group["momentum"] != 0.0 else None
            mods = opt._shadow_mods
            for b, start in zip(self._buckets, self._bucket_starts):
                for p, off in zip(b.params, b.offsets):
                    pv = torch.as_strided(new_p, p.shape, p.stride(), start + off)
                    pv.copy_(p.detach())
                    p.data = pv
                    if new_m is not None:
                        mv = torch.as_strided(new_m, p.shape, p.stride(), start + off)
                        old = opt.state[p].get("momentum_buffer")
                        if old is not None:
                            mv.copy_(old)
                        opt.state[p]["momentum_buffer"] = mv
                    m = mods.get(id(p))
                    if m is not None:        # the module's bf16 shadow becomes a view of the flat shadow array
                        m._shadow = torch.as_strided(new_s, p.shape, p.stride(), start + off)
                        m._shadow_key = None
                        m.shadow()
            self._param_region, self._shadow_region, self._mom_region = new_p, new_s, new_m
        opt._tables.clear()
        self._lr_dev = opt._lr_tensor(0, self.device, group["lr"])

    def _sgd_epilogue(self, bi: int) -> "_lib.SgdEpilogue":
        group = self._fused_opt.param_groups[0]
        start = self._bucket_starts[bi]
        e = _lib.SgdEpilogue()
        e.param = self._param_region.data_ptr() + 4 * start
        e.momentum_buf = (self._mom_region.data_ptr() + 4 * start) if self._mom_region is not None else None
        e.shadow_bf16 = self._shadow_region.data_ptr() + 2 * start
        e.lr_dev = self._lr_dev.data_ptr()
        e.momentum, e.weight_decay = float(group["momentum"]), float(group["weight_decay"])
        return e

    def consume_fused_step(self) -> bool:
        """Called by ``optimizer.step()``: ``True`` when the backward that just ran already applied the update."""
        done, self._fused_stepped = getattr(self, "_fused_stepped", False), False
        return done

    # ------------------------------------------------------------------ public surface
    @contextlib.contextmanager
    def no_sync(self):
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def register_comm_hook(self, hook) -> None:
        """Replace the gradient averaging of a bucket: ``hook(flat)`` is called once per finished bucket (on the
        communication stream on CUDA) and must leave the averaged gradients in ``flat`` (a 1-D view of the bucket).
        The analogue of torch DDP's ``register_comm_hook`` for synchronous hooks, e.g. compression or logging around
        ``ddp.backend.all_reduce_avg``.  Switches the native launch path off (the hook is Python)."""
        if getattr(self, "_fused_opt", None) is not None:
            raise RuntimeError("a communication hook cannot be combined with fuse_optimizer()")
        self._comm_hook = hook
        self._native = False

    def enable_comm_timing(self, on: bool = True) -> None:
        """Time every bucket's gradient averaging (CUDA events on the communication stream; host clock on CPU) - the
        communication-time counter next to ``utils.metrics.StepMeter``.  ``comm_times_ms()`` returns the last
        backward's numbers.  Eager steps only (timing events cannot be read out of a replayed CUDA graph); while it is
        on, buckets are launched from Python instead of by the native reducer."""
        if on and getattr(self, "_fused_opt", None) is not None:
            raise RuntimeError("communication timing cannot be combined with fuse_optimizer()")
        self._time_comm = bool(on)
        self._comm_ms = []
        self._bind_native()

    def comm_times_ms(self) -> List[float]:
        """Milliseconds the averaging of each bucket took in the last synchronised backward, in launch order
        (waits for the communication stream)."""
        return list(self._comm_ms)

    def bucket_sizes_bytes(self) -> List[int]:
        return [b.flat.numel() * b.flat.element_size() for b in self._buckets]

    def forward(self, *inputs, **kwargs):
        if self.training and torch.is_grad_enabled():
            self._prepare_iteration()
        # Buffers: rank 0's win before a forward that follows a gradient-synchronised one - torch's rule
        # (``require_forward_param_sync``), which also covers the first evaluation pass after training (ranks then
        # evaluate with the same BatchNorm statistics) and skips the broadcast between no_sync() passes.
        if self._require_forward_param_sync and self.broadcast_buffers and self.world_size > 1 and self._buf_flats:
            with torch.no_grad():
                for bf in self._buf_flats:
                    self.backend.broadcast(bf, 0)
        out = self.module(*inputs, **kwargs)
        self._require_forward_param_sync = torch.is_grad_enabled() and self.require_backward_grad_sync
        return out

    def _prepare_iteration(self) -> None:
        if self._want_rebuild and not self._rebuilt and self.stats["iterations"] >= 1:
            # one-time rebuild in the order gradients actually became ready (rank 0's order, like torch)
            box = [self._last_ready_order]
            if self.world_size > 1:
                dist.broadcast_object_list(box, src=0, group=self.process_group)
            order = box[0]
            if order is not None and len(order) == len(self._params):
                if self.device.type == "cuda":
                    torch.cuda.current_stream(self.device).synchronize()
                self._build_buckets(order)
            self._rebuilt = True
        # gradients dropped by optimizer.zero_grad(set_to_none=True): zero their bucket views and re-attach them
        # (kernels accumulate into the views); gradients that are still set keep their contents
        dropped = [p for p in self._params if p.grad is None]
        if dropped:
            if len(dropped) == len(self._params):
                self._grad_region.zero_()                # one fill for all buckets
            else:
                for p in dropped:
                    self._view_of[id(p)].zero_()
            for p in dropped:
                p.grad = self._view_of[id(p)]

    _last_ready_order: Optional[List[int]] = None

    # ------------------------------------------------------------------ backward-time machinery
    def _grad_ready(self, p: nn.Parameter) -> None:
        if not self.require_backward_grad_sync:
            return
        view = self._view_of[id(p)]
        g = p.grad
        if g is not view and (g.data_ptr() != view.data_ptr() or g.stride() != view.stride()):
            with torch.no_grad():
                view.copy_(g)
            p.grad = view
        self.mark_ready(p)

    # -- gradient sink protocol (ops.functional.set_grad_sink) --------------------------------------------------
    def targets(self, params) -> Optional[List[torch.Tensor]]:
        """Bucket views the kernels may accumulate the gradients of ``params`` into, or ``None`` to decline (a
        parameter of another module, a gradient that is not attached to its view, direct gradients switched off)."""
        if not self.direct_grads:
            return None
        out = []
        for p in params:
            v = self._view_of.get(id(p))
            g = p.grad
            if v is None or g is None or not p.requires_grad:
                return None
            if g is not v and (g.data_ptr() != v.data_ptr() or g.stride() != v.stride()):
                return None
            out.append(v)
        self.stats["direct_grads"] += len(out)
        return out

    def done(self, params) -> None:
        for p in params:
            self.mark_ready(p)

    def mark_ready(self, p: nn.Parameter) -> None:
        """Tell the reducer that ``p``'s gradient (already in its bucket view) is final for this backward."""
        if not self.require_backward_grad_sync:
            return
        if not self._callback_queued:
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
            self._callback_queued = True
        if self._native:
            # one native call: mark, and if this completed a bucket: event, stream switch, all-reduce kernel, event
            b = _lib.call_ret("ptd_reducer_mark_ready_launch", self._reducer_id, self._index_of[id(p)],
                              torch.cuda.current_stream(self.device).cuda_stream, self._comm_stream.cuda_stream)
            if b >= 0:
                bk = self._buckets[b]
                self.stats["allreduce_launches"] += 1
                self.stats["bytes_reduced"] += bk.flat.numel() * bk.flat.element_size()
                self.backend.count(self._bucket_comms[b])
                _lib.LAUNCHES += 1
            elif b < -2:
                raise RuntimeError(f"reducer error {b}: {_lib.require().ptd_last_error().decode()}")
            return
        b = _lib.call_ret("ptd_reducer_mark_ready", self._reducer_id, self._index_of[id(p)])
        if b >= 0:
            self._launch_bucket(self._buckets[b])
        elif b < -2:
            raise RuntimeError(f"reducer error {b}")

    def _launch_bucket(self, b: _Bucket) -> None:
        b.launched = True
        self.stats["allreduce_launches"] += 1
        self.stats["bytes_reduced"] += b.flat.numel() * b.flat.element_size()
        if self.world_size == 1 or self._skip_allreduce:
            return
        if self._comm_stream is not None:
            b.ready_event.record(torch.cuda.current_stream(self.device))
            self._comm_stream.wait_event(b.ready_event)
            with torch.cuda.stream(self._comm_stream), nvtx_range("ddp/bucket_allreduce"):
                if self._time_comm:
                    b.t0, b.t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    b.t0.record(self._comm_stream)
                (self._comm_hook or self.backend.all_reduce_avg)(b.flat)
                if self._time_comm:
                    b.t1.record(self._comm_stream)
                b.done_event.record(self._comm_stream)
            if self._time_comm:
                self._timed_order.append(b)
        else:
            t0 = time.perf_counter()
            (self._comm_hook or self.backend.all_reduce_avg)(b.flat)
            if self._time_comm:
                b.t0, b.t1 = t0, time.perf_counter()
                self._timed_order.append(b)

    def _finalize_backward(self) -> None:
        if self.find_unused_parameters and self.require_backward_grad_sync:
            # whoever got no gradient in this backward is reported now, in registration order (the same on every
            # rank): their views hold zeros, their buckets complete and are averaged like the others
            n = len(self._buckets)
            out = (ctypes.c_int * n)()
            if _lib.call_ret("ptd_reducer_incomplete", self._reducer_id, ctypes.cast(out, ctypes.c_void_p), n) > 0:
                for p in self._params:
                    self.mark_ready(p)              # already-marked parameters are ignored by the reducer (-2)
        self._callback_queued = False
        if self._native:
            rc = _lib.call_ret("ptd_reducer_finalize", self._reducer_id, torch.cuda.current_stream(self.device).cuda_stream)
            if rc < 0:
                raise RuntimeError(f"reducer error {rc}: {_lib.require().ptd_last_error().decode()}")
            if rc == 1 and getattr(self, "_fused_opt", None) is not None:
                self._fused_stepped = True          # every bucket's kernel also applied its SGD step
        if _lib.call_ret("ptd_reducer_all_done", self._reducer_id) != 1:
            n = len(self._buckets)
            out = (ctypes.c_int * n)()
            c = _lib.call_ret("ptd_reducer_incomplete", self._reducer_id, ctypes.cast(out, ctypes.c_void_p), n)
            _lib.call_ret("ptd_reducer_reset", self._reducer_id)
            raise RuntimeError(f"{c} gradient bucket(s) did not complete in this backward: some parameter did not "
                               "receive a gradient (pass find_unused_parameters=True)")
        if self._comm_stream is not None:
            cur = torch.cuda.current_stream(self.device)
            for b in self._buckets:
                if b.launched and self.world_size > 1 and not self._skip_allreduce:
                    cur.wait_event(b.done_event)
        for b in self._buckets:
            b.launched = False
        if self._time_comm:
            ms = []
            for b in self._timed_order:
                if self._comm_stream is not None:
                    b.t1.synchronize()
                    ms.append(b.t0.elapsed_time(b.t1))
                else:
                    ms.append((b.t1 - b.t0) * 1e3)
            self._comm_ms, self._timed_order = ms, []
        if not self._rebuilt:
            n = len(self._params)
            out = (ctypes.c_int * n)()
            got = _lib.call_ret("ptd_reducer_ready_order", self._reducer_id, ctypes.cast(out, ctypes.c_void_p), n)
            self._last_ready_order = list(out)[:got] if got == n else None
        _lib.call_ret("ptd_reducer_reset", self._reducer_id)
        self.stats["iterations"] += 1

    def __del__(self):
        try:
            if _F.grad_sink() is self:
                _F.set_grad_sink(None)
            if getattr(self, "_reducer_id", 0):
                _lib.call_ret("ptd_reducer_destroy", self._reducer_id)
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass
