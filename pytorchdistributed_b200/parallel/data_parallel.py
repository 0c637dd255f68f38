"""``DataParallel`` - single process, one replica per GPU, replicate / scatter / parallel-apply / gather.

Same surface as the ``nn.DataParallel(model)`` call of the reference
(``R/01 数据并行/01_multi_gpus_data_parallelism.ipynb#cell5-6``): ``device_ids=None`` means every visible
GPU, ``output_device=None`` means ``device_ids[0]``; ``model(input)`` takes a batch on ``device_ids[0]``
(or the host), splits it on dim 0, runs the replicas concurrently and returns the concatenated
output on the output device.

What is different underneath:

* replicas are built ONCE; each keeps its parameters and buffers as views into one flat block per
  device.  ``replicate`` is then one device-issued peer-to-peer copy kernel per GPU (peers pull the
  root's flat block over NVLink), skipped altogether while the root parameters are unchanged
  (version counters) - torch re-creates module shells and re-broadcasts every forward;
* ``scatter`` / ``gather`` are the same copy kernel: a replica pulls its input rows from the root and
  pushes its output rows straight into the root's output tensor (no concat pass);
* with ``criterion=`` the loss is computed per replica and only the per-replica scalar crosses NVLink
  ("gather fused into the loss");
* backward runs each replica's own autograd graph on its device (threads), gradients land in flat
  per-device blocks and one reduce-to-root kernel sums them into the root's ``.grad`` (replaces
  ``ReduceAddCoalesced``).
"""
from __future__ import annotations

import copy
import ctypes
import threading
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, List, Optional, Sequence

import torch
from torch import nn

from .. import _lib
from ..comm import p2p as _p2p  # noqa: F401  (registers the entry points)


def _peer_copy(dst: torch.Tensor, src: torch.Tensor, nbytes: int, device: torch.device) -> None:
    """Copy kernel running on ``device`` (current stream of that device); src/dst may live on any peer."""
    with torch.cuda.device(device):
        _lib.call("ptd_peer_copy", dst, src, nbytes, torch.cuda.current_stream(device).cuda_stream)


def _dense_flat(t: torch.Tensor) -> torch.Tensor:
    return torch.as_strided(t, (t.numel(),), (1,), t.storage_offset())


class _FlatBlock:
    """All tensors of one kind (params / grads / buffers of a dtype) of a module as views of one buffer."""

    def __init__(self, tensors: List[torch.Tensor], device: torch.device, dtype: torch.dtype) -> None:
        self.offsets, total = [], 0
        for t in tensors:
            total = (total + 3) // 4 * 4
            self.offsets.append(total)
            total += t.numel()
        self.total = max((total + 3) // 4 * 4, 4)
        self.flat = torch.zeros(self.total, device=device, dtype=dtype)

    def view_like(self, i: int, t: torch.Tensor) -> torch.Tensor:
        return torch.as_strided(self.flat, t.shape, t.stride(), self.offsets[i])


class _Replica:
    def __init__(self, module: nn.Module, device: torch.device, is_root: bool) -> None:
        self.device = device
        self.module = module
        self.params = [p for p in module.parameters()]
        with torch.no_grad():
            self.pblock = _FlatBlock(self.params, device, torch.float32)
            for i, p in enumerate(self.params):
                v = self.pblock.view_like(i, p)
                v.copy_(p.detach())
                p.data = v
            self.gblock = _FlatBlock(self.params, device, torch.float32)
            self.gviews = [self.gblock.view_like(i, p) for i, p in enumerate(self.params)]
            self.fbufs = [b for b in module.buffers() if b.is_floating_point()]
            self.ibufs = [b for b in module.buffers() if not b.is_floating_point()]
            self.bblock = _FlatBlock(self.fbufs, device, torch.float32) if self.fbufs else None
            if self.bblock is not None:
                owners = {id(b): (m, n) for m in module.modules() for n, b in m._buffers.items() if b is not None}
                for i, b in enumerate(self.fbufs):
                    v = self.bblock.view_like(i, b)
                    v.copy_(b)
                    m, n = owners[id(b)]
                    m._buffers[n] = v
        self.stream = torch.cuda.current_stream(device)

    def attach_grads(self) -> None:
        for p, g in zip(self.params, self.gviews):
            p.grad = g


class DataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids: Optional[Sequence[int]] = None, output_device=None,
                 dim: int = 0, criterion: Optional[Callable] = None) -> None:
        super().__init__()
        if not torch.cuda.is_available():
            self.module, self.device_ids, self.criterion = module, [], criterion
            return
        if dim != 0:
            raise NotImplementedError("DataParallel splits on dim 0")
        if device_ids is None:
            device_ids = list(range(torch.cuda.device_count()))
        self.device_ids = [d if isinstance(d, int) else torch.device(d).index for d in device_ids]
        self.output_device = self.device_ids[0] if output_device is None else (
            output_device if isinstance(output_device, int) else torch.device(output_device).index)
        self.criterion = criterion
        self.module = module
        root = torch.device("cuda", self.device_ids[0])
        first = next(module.parameters(), None)
        if first is not None and first.device.type != "cuda":
            module.to(root)             # torch only auto-moves for a single device; we do it always
        elif first is not None and first.device != root:
            raise RuntimeError(f"module must have its parameters on device {root} (device_ids[0]) but found {first.device}")
        for a in self.device_ids:
            for b in self.device_ids:
                if a != b:
                    _lib.call("ptd_p2p_enable_peer_access", a, b)
        self.replicas: List[_Replica] = [_Replica(module, root, True)]
        for d in self.device_ids[1:]:
            dev = torch.device("cuda", d)
            with torch.cuda.device(dev):
                self.replicas.append(_Replica(copy.deepcopy(module).to(dev), dev, False))
        self._pool = ThreadPoolExecutor(max_workers=max(1, len(self.device_ids)))
        self._param_version = None
        self._anchor = torch.zeros(1, device=root, requires_grad=True)   # makes the fused node part of the graph
        self.stats = {"replicate_copies": 0, "replicate_skipped": 0}
        self._lock = threading.Lock()

    # -------------------------------------------------------------------------------- replicate
    def _versions(self):
        return tuple(p._version for p in self.replicas[0].params)

    def replicate(self, force: bool = False) -> None:
        """Bring every replica's parameters/buffers in line with the root (one pull kernel per GPU)."""
        root = self.replicas[0]
        ver = self._versions()
        need_params = force or ver != self._param_version
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(root.device))
        for rep in self.replicas[1:]:
            with torch.cuda.device(rep.device):
                st = torch.cuda.current_stream(rep.device)
                st.wait_event(ev)
                if need_params:
                    _peer_copy(rep.pblock.flat, root.pblock.flat, root.pblock.total * 4, rep.device)
                    # the raw kernel does not touch version counters: bump them so that everything cached on
                    # (data_ptr, version) - the bf16 weight shadows of Conv2d / Linear - is rebuilt from the new values
                    torch.autograd.graph.increment_version(rep.params)
                    self.stats["replicate_copies"] += 1
                if rep.bblock is not None:
                    _peer_copy(rep.bblock.flat, root.bblock.flat, root.bblock.total * 4, rep.device)
                for rb, sb in zip(rep.ibufs, root.ibufs):
                    rb.copy_(sb, non_blocking=True)
        if not need_params:
            self.stats["replicate_skipped"] += 1
        self._param_version = ver

    # -------------------------------------------------------------------------------- forward
    def _chunks(self, n: int) -> List[int]:
        k = len(self.device_ids)
        per = (n + k - 1) // k
        sizes = [max(0, min(per, n - i * per)) for i in range(k)]
        return [s for s in sizes if s > 0]

    def _scatter(self, t: torch.Tensor, sizes: List[int]) -> List[torch.Tensor]:
        """Row slices of ``t``, one per replica, each on its replica's device."""
        outs, start = [], 0
        t = t if (not t.is_cuda) else t.contiguous()
        row_bytes = (t.numel() // max(t.size(0), 1)) * t.element_size()
        ev = None
        if t.is_cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(t.device))
        for rep, sz in zip(self.replicas, sizes):
            sl = t[start:start + sz]
            if t.is_cuda and t.device == rep.device:
                outs.append(sl)
            elif not t.is_cuda:
                outs.append(sl.to(rep.device, non_blocking=True))
            else:
                with torch.cuda.device(rep.device):
                    dst = torch.empty(sl.shape, device=rep.device, dtype=t.dtype)
                    torch.cuda.current_stream(rep.device).wait_event(ev)
                    nb = sz * row_bytes
                    if nb % 16 == 0 and sl.data_ptr() % 16 == 0:
                        _peer_copy(dst, sl, nb, rep.device)          # the replica pulls its rows
                    else:
                        dst.copy_(sl, non_blocking=True)
                outs.append(dst)
            start += sz
        return outs

    def forward(self, *inputs, **kwargs):
        if not self.device_ids:
            out = self.module(*inputs, **kwargs)
            return out
        if len(self.device_ids) == 1:
            return self.module(*inputs, **kwargs)
        x = inputs[0]
        sizes = self._chunks(x.size(0))
        xs = self._scatter(x, sizes)
        extra = [self._scatter(a, sizes) if isinstance(a, torch.Tensor) and a.dim() > 0 and a.size(0) == x.size(0)
                 else [a] * len(sizes) for a in inputs[1:]]
        self.replicate()
        grad_mode = torch.is_grad_enabled()
        return _ParallelApply.apply(self, grad_mode, sizes, x.size(0), xs, extra, kwargs, self._anchor)

    def parallel_apply(self, fn_per_replica: Callable[[int], object], n: int) -> List:
        futs = [self._pool.submit(fn_per_replica, i) for i in range(n)]
        return [f.result() for f in futs]


class _ParallelApply(torch.autograd.Function):
    """Runs the replicas (each with its own autograd graph), gathers outputs on the output device; the
    backward scatters the incoming gradient, runs the replica backward passes and reduces to the root."""

    @staticmethod
    def forward(ctx, dp: DataParallel, grad_mode: bool, sizes, total, xs, extra, kwargs, _anchor):
        reps = dp.replicas[:len(sizes)]
        out_dev = torch.device("cuda", dp.output_device)
        criterion = dp.criterion

        def run(i):
            rep = reps[i]
            with torch.cuda.device(rep.device), torch.set_grad_enabled(grad_mode):
                args = [e[i] for e in extra]
                if criterion is not None:
                    out = rep.module(xs[i], **kwargs)
                    loss = criterion(out, *args)
                    return loss * (sizes[i] / total)       # weight so the sum is the global mean
                return rep.module(xs[i], *args, **kwargs)

        outs = dp.parallel_apply(run, len(reps))
        ctx.dp, ctx.outs, ctx.sizes, ctx.reps = dp, outs, sizes, reps
        ctx.fused_loss = criterion is not None
        with torch.no_grad():
            if criterion is not None:
                slots = torch.zeros(len(reps), 4, device=out_dev, dtype=torch.float32)    # 16 B per replica
                evs = []
                for i, (rep, o) in enumerate(zip(reps, outs)):
                    with torch.cuda.device(rep.device):
                        src = torch.zeros(4, device=rep.device, dtype=torch.float32)
                        src[0] = o.detach().float()
                        _peer_copy(slots[i], src, 16, rep.device)      # only the scalar crosses NVLink
                        ev = torch.cuda.Event()
                        ev.record(torch.cuda.current_stream(rep.device))
                        evs.append((ev, src))
                for ev, _ in evs:
                    torch.cuda.current_stream(out_dev).wait_event(ev)
                ctx.keep = evs
                return slots[:, 0].sum()
            o0 = outs[0]
            result = torch.empty((total,) + tuple(o0.shape[1:]), device=out_dev, dtype=o0.dtype)
            row_bytes = (o0.numel() // max(o0.size(0), 1)) * o0.element_size()
            start, evs = 0, []
            for rep, o, sz in zip(reps, outs, sizes):
                dst = result[start:start + sz]
                with torch.cuda.device(rep.device):
                    oc = o.detach().contiguous()
                    nb = sz * row_bytes
                    if nb % 16 == 0 and dst.data_ptr() % 16 == 0 and oc.data_ptr() % 16 == 0:
                        _peer_copy(dst, oc, nb, rep.device)            # the replica pushes its rows into the root tensor
                    else:
                        dst.copy_(oc, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(torch.cuda.current_stream(rep.device))
                    evs.append((ev, oc))
                start += sz
            for ev, _ in evs:
                torch.cuda.current_stream(out_dev).wait_event(ev)
            ctx.keep = evs
            return result

    @staticmethod
    def backward(ctx, grad_out):
        dp, outs, sizes, reps = ctx.dp, ctx.outs, ctx.sizes, ctx.reps
        root = reps[0]
        ev0 = torch.cuda.Event()
        ev0.record(torch.cuda.current_stream(grad_out.device))
        if ctx.fused_loss:
            gs = [grad_out.to(rep.device, non_blocking=True) for rep in reps]
        else:
            gs, start = [], 0
            go = grad_out.contiguous()
            for rep, sz in zip(reps, sizes):
                sl = go[start:start + sz]
                gs.append(sl if sl.device == rep.device else sl.to(rep.device, non_blocking=True))
                start += sz

        # peers start from zero every backward; the root keeps whatever the user already accumulated
        old = [p.grad for p in root.params]
        with torch.no_grad():
            if all(g is None for g in old):
                root.gblock.flat.zero_()
            elif not all(g is not None and g.data_ptr() == v.data_ptr() for g, v in zip(old, root.gviews)):
                for g, v in zip(old, root.gviews):
                    if g is None:
                        v.zero_()
                    elif g.data_ptr() != v.data_ptr():
                        v.copy_(g)
            for rep in reps[1:]:
                with torch.cuda.device(rep.device):
                    rep.gblock.flat.zero_()
        for rep in reps:
            rep.attach_grads()

        def run(i):
            rep = reps[i]
            with torch.cuda.device(rep.device):
                torch.cuda.current_stream(rep.device).wait_event(ev0)
                if outs[i].requires_grad:
                    torch.autograd.backward(outs[i], gs[i].to(outs[i].dtype).reshape(outs[i].shape))
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(rep.device))
                return ev

        # The replica that lives on this node's own device must run its backward on THIS thread (re-entrant
        # autograd): its nodes belong to the engine thread we are occupying, a pool thread would wait for it forever.
        # The other replicas run from pool threads and overlap with it.
        here = grad_out.device
        futs = {i: dp._pool.submit(run, i) for i in range(len(reps)) if reps[i].device != here}
        evs = []
        for i in range(len(reps)):
            if reps[i].device == here:
                evs.append(run(i))
        evs += [f.result() for f in futs.values()]
        with torch.cuda.device(root.device), torch.no_grad():
            st = torch.cuda.current_stream(root.device)
            for ev in evs:
                st.wait_event(ev)
            srcs = (ctypes.c_void_p * (len(reps) - 1))(*[r.gblock.flat.data_ptr() for r in reps[1:]])
            # root.grad_flat = root.grad_flat + sum(peer grad_flat): the root reads its peers over NVLink
            _lib.call("ptd_reduce_sum_f32", root.gblock.flat, ctypes.cast(srcs, ctypes.c_void_p), len(reps) - 1,
                      root.gblock.total, 1.0, 1, st.cuda_stream)
        return (None,) * 8
