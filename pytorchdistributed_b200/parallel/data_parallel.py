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
  ``ReduceAddCoalesced``);
* weights the tensor-core kernels read as bf16 travel as bf16: the root keeps the bf16 shadows of all its
  convolution / classifier weights in one flat block and a replica pulls THAT block (half the bytes of the fp32
  masters, and no cast kernels on the replica); only the small fp32 parameters (BatchNorm, biases, the stem) are
  pulled as fp32;
* ``cuda_graph=True`` (training mode, fixed batch shape): every replica's forward and backward are captured once
  as a CUDA-graph pair and replayed from ONE thread - a single Python process cannot issue ~700 kernel launches
  per replica and step fast enough for 4-8 B200s (the eager path ran 4 GPUs slower than one), a graph launch
  per replica and direction it can.  Weight gradients land in the flat per-device blocks directly (gradient sink
  of ``ops.functional``), the loss scalar / output rows are pushed to the root inside the forward graph.
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
        al = 16 // torch.empty((), dtype=dtype).element_size()       # every view starts 16-byte aligned
        for t in tensors:
            total = (total + al - 1) // al * al
            self.offsets.append(total)
            total += t.numel()
        self.total = max((total + al - 1) // al * al, al)
        self.esize = 16 // al
        self.flat = torch.zeros(self.total, device=device, dtype=dtype)

    def view_like(self, i: int, t: torch.Tensor) -> torch.Tensor:
        return torch.as_strided(self.flat, t.shape, t.stride(), self.offsets[i])


def _shadowed(module: nn.Module):
    """(module, weight) pairs whose forward reads a bf16 shadow of the weight that is a plain cast of it (Conv2d but
    not the stem, whose shadow is a re-packed tensor; Linear but not the tutorial-sized fp32 one)."""
    out = []
    for m in module.modules():
        if hasattr(m, "shadow") and hasattr(m, "mark_shadow_fresh") and not getattr(m, "is_stem", False) \
                and not getattr(m, "small", False) and isinstance(getattr(m, "weight", None), nn.Parameter):
            out.append(m)
    return out


class _ReplicaSink:
    """Gradient sink (``ops.functional.set_grad_sink``) of one replica while its backward graph is captured: the
    weight-gradient kernels accumulate straight into the replica's flat gradient block."""

    def __init__(self, params, views) -> None:
        self.map = {id(p): v for p, v in zip(params, views)}

    def targets(self, params):
        out = [self.map.get(id(p)) for p in params]
        return None if any(v is None for v in out) else out

    def done(self, params) -> None:
        pass


class _Replica:
    def __init__(self, module: nn.Module, device: torch.device, is_root: bool) -> None:
        self.device = device
        self.module = module
        # parameters WITHOUT a bf16 shadow first: a replica refreshes that fp32 prefix and the shadow block, not the
        # fp32 masters of the big weights (nothing on a replica reads them)
        self.shadow_mods = _shadowed(module)
        big = {id(m.weight) for m in self.shadow_mods}
        allp = [p for p in module.parameters()]
        self.params = [p for p in allp if id(p) not in big] + [p for p in allp if id(p) in big]
        self.n_small = len(allp) - len(big)
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
            owners = {id(b): (m, n) for m in module.modules() for n, b in m._buffers.items() if b is not None}
            if self.bblock is not None:
                for i, b in enumerate(self.fbufs):
                    v = self.bblock.view_like(i, b)
                    v.copy_(b)
                    m, n = owners[id(b)]
                    m._buffers[n] = v
            # integer buffers (BatchNorm's num_batches_tracked) in one block too: one copy kernel per replica instead
            # of one cross-device torch copy per buffer (each of which fences both devices' streams)
            i64 = [b for b in self.ibufs if b.dtype == torch.int64]
            self.iblock = _FlatBlock(i64, device, torch.int64) if (i64 and device.type == "cuda") else None
            if self.iblock is not None:
                for i, b in enumerate(i64):
                    v = self.iblock.view_like(i, b)
                    v.copy_(b)
                    m, n = owners[id(b)]
                    m._buffers[n] = v
                self.ibufs = [b for b in self.ibufs if b.dtype != torch.int64]
            # fp32 elements of the "small" prefix (what a replica pulls as fp32)
            self.small_total = self.pblock.offsets[self.n_small] if self.n_small < len(self.params) else self.pblock.total
            # bf16 shadows of the big weights as views of one flat block (same layout on every replica)
            self.sblock = None
            if self.shadow_mods and device.type == "cuda":
                ws = [m.weight for m in self.shadow_mods]
                self.sblock = _FlatBlock(ws, device, torch.bfloat16)
                for i, m in enumerate(self.shadow_mods):
                    m._shadow = self.sblock.view_like(i, m.weight)
                    m._shadow_key = None
        self.stream = torch.cuda.current_stream(device)

    def refresh_shadows(self) -> None:
        """Root only: bring the bf16 shadows up to date (no launch for those an attached fused optimizer refreshed)."""
        for i, m in enumerate(self.shadow_mods):
            view = self.sblock.view_like(i, m.weight)
            if m._shadow is None or m._shadow.data_ptr() != view.data_ptr():      # dropped by a later .to() / _apply
                m._shadow, m._shadow_key = view, None
            m.shadow()

    def shadows_pulled(self) -> None:
        """Replica only: the shadow block was just overwritten with the root's; the cached casts are current."""
        for i, m in enumerate(self.shadow_mods):
            view = self.sblock.view_like(i, m.weight)
            if m._shadow is None or m._shadow.data_ptr() != view.data_ptr():
                m._shadow = view
            m.mark_shadow_fresh()

    def attach_grads(self) -> None:
        for p, g in zip(self.params, self.gviews):
            p.grad = g


class DataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids: Optional[Sequence[int]] = None, output_device=None,
                 dim: int = 0, criterion: Optional[Callable] = None, cuda_graph: bool = False) -> None:
        super().__init__()
        self.cuda_graph = bool(cuda_graph)
        self._graphed = None            # (key, [_GraphedReplica]) once captured
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
        self.stats = {"replicate_copies": 0, "replicate_skipped": 0, "graph_replays": 0, "graph_captures": 0}
        self._lock = threading.Lock()

    def train(self, mode: bool = True):
        super().train(mode)
        for rep in getattr(self, "replicas", [])[1:]:       # replicas are not sub-modules: keep their mode in step
            rep.module.train(mode)
        return self

    # -------------------------------------------------------------------------------- replicate
    def _versions(self):
        return tuple(p._version for p in self.replicas[0].params)

    def replicate(self, force: bool = False) -> None:
        """Bring every replica's parameters/buffers in line with the root (one pull kernel per GPU)."""
        root = self.replicas[0]
        ver = self._versions()
        need_params = force or ver != self._param_version
        if need_params and root.sblock is not None:
            with torch.cuda.device(root.device):
                root.refresh_shadows()
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(root.device))
        for rep in self.replicas[1:]:
            with torch.cuda.device(rep.device):
                st = torch.cuda.current_stream(rep.device)
                st.wait_event(ev)
                if need_params:
                    if root.sblock is not None:
                        # fp32 prefix (BatchNorm, biases, stem, small layers) + the bf16 shadows of everything else
                        _peer_copy(rep.pblock.flat, root.pblock.flat, root.small_total * 4, rep.device)
                        _peer_copy(rep.sblock.flat, root.sblock.flat, root.sblock.total * 2, rep.device)
                    else:
                        _peer_copy(rep.pblock.flat, root.pblock.flat, root.pblock.total * 4, rep.device)
                    # the raw kernel does not touch version counters: bump them so that everything cached on
                    # (data_ptr, version) - e.g. the packed stem weights - is rebuilt from the new values ...
                    torch.autograd.graph.increment_version(rep.params)
                    rep.shadows_pulled()             # ... except the shadows that just arrived ready-made
                    self.stats["replicate_copies"] += 1
                if rep.bblock is not None:
                    _peer_copy(rep.bblock.flat, root.bblock.flat, root.bblock.total * 4, rep.device)
                if rep.iblock is not None:
                    _peer_copy(rep.iblock.flat, root.iblock.flat, root.iblock.total * 8, rep.device)
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
            if self.criterion is not None:
                return self.criterion(self.module(inputs[0], **kwargs), *inputs[1:])
            return self.module(*inputs, **kwargs)
        x = inputs[0]
        if self.cuda_graph and not kwargs and self.module.training and torch.is_grad_enabled():
            return self._forward_graphed(inputs)
        sizes = self._chunks(x.size(0))
        xs = self._scatter(x, sizes)
        extra = [self._scatter(a, sizes) if isinstance(a, torch.Tensor) and a.dim() > 0 and a.size(0) == x.size(0)
                 else [a] * len(sizes) for a in inputs[1:]]
        self.replicate()
        grad_mode = torch.is_grad_enabled()
        return _ParallelApply.apply(self, grad_mode, sizes, x.size(0), xs, extra, kwargs, self._anchor)

    # -------------------------------------------------------------------------------- CUDA-graph mode
    def _graph_key(self, inputs):
        return tuple((tuple(a.shape), a.dtype, a.device.type) if isinstance(a, torch.Tensor) else ("const", repr(a))
                     for a in inputs)

    def _forward_graphed(self, inputs):
        x = inputs[0]
        total = x.size(0)
        sizes = self._chunks(total)
        key = self._graph_key(inputs)
        self.replicate()
        if self._graphed is None or self._graphed[0] != key:
            self._graphed = None                 # drop the old graphs (and their memory pools) first
            self._capture(key, inputs, sizes)
        return _GraphedApply.apply(self, inputs, sizes, self._anchor)

    def _capture(self, key, inputs, sizes) -> None:
        x = inputs[0]
        total = x.size(0)
        root_dev = self.replicas[0].device
        n = len(sizes)
        self._slots = torch.zeros(n, 4, device=root_dev, dtype=torch.float32)       # 16 B per replica
        self._g_result = None
        graphed, start = [], 0
        for i, (rep, sz) in enumerate(zip(self.replicas, sizes)):
            chunk = [a[start:start + sz] if (isinstance(a, torch.Tensor) and a.dim() > 0 and a.size(0) == total) else a
                     for a in inputs]
            gr = _GraphedReplica(self, i, rep, chunk, sz / total, (start, sz, total))
            graphed.append(gr)
            start += sz
        self._graphed = (key, graphed)
        self.stats["graph_captures"] += 1

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


def _pull_rows(dst: torch.Tensor, src: torch.Tensor, device: torch.device) -> None:
    """``dst`` <- ``src`` by a copy kernel on ``device``'s current stream; either side may live on a peer GPU (a pull
    or a push over NVLink), ``src`` may be host memory (then a plain async copy)."""
    nb = src.numel() * src.element_size()
    if src.is_cuda and dst.is_cuda and nb % 16 == 0 and src.data_ptr() % 16 == 0 and dst.data_ptr() % 16 == 0 \
            and src.is_contiguous() and dst.is_contiguous():
        _peer_copy(dst, src, nb, device)
    else:
        with torch.cuda.device(device):
            dst.copy_(src, non_blocking=True)


class _GraphedReplica:
    """One replica's (forward, backward) CUDA-graph pair for a fixed input shape.

    forward graph : static inputs -> module (-> criterion * share of the batch) -> the loss scalar / the output rows
                    pushed into the root's slot / result tensor by a peer-copy kernel inside the graph;
    backward graph: static output gradient -> zero the replica's flat gradient block -> autograd with the block's views
                    as gradient sink (weight-gradient kernels accumulate there directly) -> the few gradients autograd
                    still returns are added to their views."""

    def __init__(self, dp: "DataParallel", index: int, rep: _Replica, chunk, weight: float, rows) -> None:
        from ..ops import functional as F_
        self.rep, self.index = rep, index
        dev = rep.device
        criterion = dp.criterion
        start, sz, total = rows
        params = [p for p in rep.params if p.requires_grad]
        with torch.cuda.device(dev), torch.no_grad():
            self.static_in = [torch.empty(a.shape, dtype=a.dtype, device=dev).copy_(a) if isinstance(a, torch.Tensor) else a
                              for a in chunk]
            # the root's own .grad block holds what the user accumulated: its fresh gradients go to a second block
            if index == 0:
                if not hasattr(rep, "gstage"):
                    rep.gstage = _FlatBlock(rep.params, dev, torch.float32)
                self.gblock = rep.gstage
            else:
                self.gblock = rep.gblock
            self.gviews = [self.gblock.view_like(i, p) for i, p in enumerate(rep.params)]
            buffers = [b.detach().clone() for b in rep.module.buffers()]

        def fn():
            xin = self.static_in[0]
            rest = self.static_in[1:]
            if criterion is not None:
                return criterion(rep.module(xin), *rest) * weight
            return rep.module(xin, *rest)

        with torch.cuda.device(dev):
            cur = torch.cuda.current_stream(dev)
            side = torch.cuda.Stream(dev)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                for _ in range(2):          # lazy initialisation (kernel attributes, side streams) outside any capture
                    out = fn()
                    torch.autograd.grad(out, params, torch.ones_like(out), allow_unused=True)
            cur.wait_stream(side)
            torch.cuda.synchronize(dev)
            with torch.no_grad():           # the warm-up advanced BatchNorm statistics: put them back
                for b, keep in zip(rep.module.buffers(), buffers):
                    b.copy_(keep)
            # the stem's packed weights are rebuilt by a host-side version check in eager mode; a replayed graph has
            # no host side, so it re-packs them itself every time
            for m in rep.module.modules():
                if getattr(m, "is_stem", False):
                    m._shadow_key = None
            pool = torch.cuda.graph_pool_handle()
            self.fwd = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.fwd, pool=pool, stream=side):
                out = fn()
                with torch.no_grad():
                    if criterion is not None:
                        src = torch.zeros(4, device=dev, dtype=torch.float32)
                        src[0] = out.detach().float()
                        _peer_copy(dp._slots[index], src, 16, dev)
                    else:
                        oc = out.detach().contiguous()
                        if dp._g_result is None:
                            with torch.cuda.device(dp.replicas[0].device):
                                dp._g_result = torch.empty((total,) + tuple(oc.shape[1:]), dtype=oc.dtype,
                                                           device=torch.device("cuda", dp.output_device))
                        _pull_rows(dp._g_result[start:start + sz], oc, dev)      # the replica pushes its rows
            self.out = out                                  # keeps the tensors its backward needs alive in the pool
            if criterion is not None:
                # the loss gradient (one float on the root) arrives by a 16-byte pull kernel, not by a cross-device
                # torch copy: that would make the root's stream and this one wait for each other, i.e. serialise the
                # replicas' backward graphs
                self.gout_buf = torch.zeros(4, device=dev, dtype=torch.float32)
                self.static_gout = self.gout_buf[0].view(out.shape)
            else:
                self.static_gout = torch.zeros_like(out.detach())
            sink = _ReplicaSink(rep.params, self.gviews)
            prev = F_.grad_sink()
            F_.set_grad_sink(sink)
            try:
                self.bwd = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.bwd, pool=pool, stream=side):
                    self.gblock.flat.zero_()
                    grads = torch.autograd.grad(out, params, self.static_gout, allow_unused=True)
                    with torch.no_grad():
                        by_id = {id(p): v for p, v in zip(rep.params, self.gviews)}
                        live = [(by_id[id(p)], g) for p, g in zip(params, grads) if g is not None]
                        if live:
                            torch._foreach_add_([v for v, _ in live], [g for _, g in live])
            finally:
                F_.set_grad_sink(prev)
            torch.cuda.synchronize(dev)


class _GraphedApply(torch.autograd.Function):
    """Forward / backward of all replicas as graph replays issued from the calling thread."""

    @staticmethod
    def forward(ctx, dp: DataParallel, inputs, sizes, _anchor):
        graphed = dp._graphed[1]
        x = inputs[0]
        total = x.size(0)
        root_dev = dp.replicas[0].device
        srcs = [a.contiguous() if isinstance(a, torch.Tensor) else a for a in inputs]
        ev_in = None
        if any(isinstance(a, torch.Tensor) and a.is_cuda for a in srcs):
            ev_in = torch.cuda.Event()
            ev_in.record(torch.cuda.current_stream(root_dev))
        evs, start = [], 0
        for gr, sz in zip(graphed, sizes):
            dev = gr.rep.device
            with torch.cuda.device(dev):
                st = torch.cuda.current_stream(dev)
                if ev_in is not None:
                    st.wait_event(ev_in)
                for dst, a in zip(gr.static_in, srcs):
                    if isinstance(a, torch.Tensor) and isinstance(dst, torch.Tensor):
                        rows = a[start:start + sz] if (a.dim() > 0 and a.size(0) == total) else a
                        _pull_rows(dst, rows, dev)
                gr.fwd.replay()
                ev = torch.cuda.Event()
                ev.record(st)
                evs.append(ev)
            start += sz
        out_st = torch.cuda.current_stream(torch.device("cuda", dp.output_device))
        for ev in evs:
            out_st.wait_event(ev)
        dp.stats["graph_replays"] += 1
        ctx.dp, ctx.sizes = dp, sizes
        ctx.fused_loss = dp.criterion is not None
        with torch.cuda.device(root_dev):
            if ctx.fused_loss:
                return dp._slots[:len(graphed), 0].sum()
            return dp._g_result.clone()

    @staticmethod
    def backward(ctx, grad_out):
        dp, sizes = ctx.dp, ctx.sizes
        graphed = dp._graphed[1]
        root = dp.replicas[0]
        go = grad_out.contiguous()
        if ctx.fused_loss:
            with torch.cuda.device(go.device):
                if getattr(dp, "_gout_root", None) is None or dp._gout_root.device != go.device:
                    dp._gout_root = torch.zeros(4, device=go.device, dtype=torch.float32)
                dp._gout_root[0:1].copy_(go.reshape(1).float())
        ev0 = torch.cuda.Event()
        ev0.record(torch.cuda.current_stream(go.device))
        evs, start = [], 0
        for gr, sz in zip(graphed, sizes):
            dev = gr.rep.device
            with torch.cuda.device(dev):
                st = torch.cuda.current_stream(dev)
                st.wait_event(ev0)
                if ctx.fused_loss:
                    _peer_copy(gr.gout_buf, dp._gout_root, 16, dev)
                else:
                    _pull_rows(gr.static_gout, go[start:start + sz].reshape(gr.static_gout.shape), dev)
                gr.bwd.replay()
                ev = torch.cuda.Event()
                ev.record(st)
                evs.append(ev)
            start += sz
        # root .grad = (what the user already accumulated) + sum over replicas, one kernel on the root that reads the
        # peers' flat blocks over NVLink
        old = [p.grad for p in root.params]
        with torch.cuda.device(root.device), torch.no_grad():
            st = torch.cuda.current_stream(root.device)
            accumulate = 1
            if all(g is None for g in old):
                accumulate = 0
            elif not all(g is not None and g.data_ptr() == v.data_ptr() for g, v in zip(old, root.gviews)):
                for g, v in zip(old, root.gviews):
                    if g is None:
                        v.zero_()
                    elif g.data_ptr() != v.data_ptr():
                        v.copy_(g)
            for ev in evs:
                st.wait_event(ev)
            ptrs = [gr.gblock.flat.data_ptr() for gr in graphed]
            arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
            _lib.call("ptd_reduce_sum_f32", root.gblock.flat, ctypes.cast(arr, ctypes.c_void_p), len(ptrs),
                      root.gblock.total, 1.0, accumulate, st.cuda_stream)
        root.attach_grads()
        return None, None, None, None
