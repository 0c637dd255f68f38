"""CUDA-graph replay of a pipeline stage, one (forward, backward) graph pair per micro-batch in flight.

A micro-batch pipeline is the textbook launch-bound loop: at ``split_size=1`` the reference's
``PipelineParallelResNet50`` (``R/03 模型并行/03_model_parallel.ipynb#cell12-13``) spends 4.87 s per three
iterations issuing ~100 small kernels per micro-batch and stage from Python.  Here every micro-batch of a stage
replays a captured graph instead - one ``cudaGraphLaunch`` for its forward, one for its backward.

GPipe order (all forwards, then all backwards) keeps the activations of every micro-batch alive at once, so a
stage needs one graph pair *per micro-batch*, each with its own saved tensors; they share one memory pool and use
the same memory an eager run would.  Parameter gradients come out of each backward graph in static buffers and
are added to ``.grad`` with one ``torch._foreach_add_`` per micro-batch (not one ``AccumulateGrad`` per
parameter).

bf16 weight shadows (``Conv2d.shadow()``) are refreshed by host-side version checks in eager mode; a replayed
graph has no host side, so the FIRST micro-batch's forward graph is captured with every shadow marked stale and
therefore re-casts all weights each step before anything uses them.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
from torch import nn


def _dense(t: torch.Tensor) -> bool:
    return t.is_contiguous() or (t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last))


def _mark_shadows_stale(module: nn.Module) -> None:
    for m in module.modules():
        if hasattr(m, "_shadow_key"):
            m._shadow_key = None


def _order_after(consumer: torch.device, producer: torch.device) -> None:
    """``consumer``'s current stream waits for everything issued so far on ``producer``'s current stream."""
    if consumer != producer:
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(producer))
        torch.cuda.current_stream(consumer).wait_event(ev)


class _Replay(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, anchor: torch.Tensor, owner: "GraphedStage", i: int):
        ctx.owner, ctx.i = owner, i
        with torch.cuda.device(owner.device):           # replay launches on the current device's stream
            if owner.pull_from is not None and x.data_ptr() == owner.pull_from[i].data_ptr():
                # the forward graph itself pulls its input from the producer stage's static output (peer memory):
                # only the ordering between the two devices' streams is left to do here
                _order_after(owner.device, x.device)
            else:
                owner.static_in[i].copy_(x)
                if owner.pull_from is not None:      # unexpected input: the captured pull would overwrite it
                    raise RuntimeError("graphed stage with a captured input pull got a different input tensor")
            owner.fwd[i].replay()
        return owner.static_out[i].detach()

    @staticmethod
    def backward(ctx, g: torch.Tensor):
        owner, i = ctx.owner, ctx.i
        with torch.cuda.device(owner.device):
            if g.data_ptr() != owner.static_gout[i].data_ptr():      # (a consumer stage may have pushed it there already)
                owner.static_gout[i].copy_(g)
            owner.bwd[i].replay()
            owner.accumulate(i)
        if owner.push_gin_to is not None:
            # the backward graph pushed the input gradient into the producer stage's buffer on the other device
            gin = owner.push_gin_to[i]
            _order_after(gin.device, owner.device)
            return gin, None, None, None
        gin = owner.static_gin[i]
        return gin, None, None, None


class GraphedStage:
    """``n`` graph pairs for ``module`` on ``device`` with inputs shaped like ``sample`` (one micro-batch)."""

    def __init__(self, module: nn.Module, device: torch.device, sample: torch.Tensor, n: int,
                 input_requires_grad: bool, pull_from: Optional[Sequence[torch.Tensor]] = None,
                 push_gin_to: Optional[Sequence[torch.Tensor]] = None,
                 out_slices: Optional[Sequence[torch.Tensor]] = None) -> None:
        """``pull_from[i]``: tensor (usually on the previous stage's GPU) that micro-batch i's forward graph copies
        into its static input with a captured peer-copy kernel; ``push_gin_to[i]``: tensor (on that GPU) its backward
        graph writes the input gradient into - the stage boundary then costs no launch of its own.  ``out_slices[i]``:
        preallocated block handed to the module as ``out=`` (the classifier writes its logits there)."""
        from .. import _lib
        self.module, self.device, self.n = module, device, n
        self.pull_from = list(pull_from) if pull_from is not None else None
        self.push_gin_to = list(push_gin_to) if push_gin_to is not None else None

        def peer_copy(dst: torch.Tensor, src: torch.Tensor) -> None:
            nb = src.numel() * src.element_size()
            if nb % 16 == 0 and dst.data_ptr() % 16 == 0 and src.data_ptr() % 16 == 0:
                _lib.call("ptd_peer_copy", dst, src, nb, torch.cuda.current_stream(device).cuda_stream)
            else:
                dst.copy_(src, non_blocking=True)

        self.params: List[nn.Parameter] = [p for p in module.parameters() if p.requires_grad]
        self.fwd: List[torch.cuda.CUDAGraph] = []
        self.bwd: List[torch.cuda.CUDAGraph] = []
        self.static_in: List[torch.Tensor] = []
        self.static_out: List[torch.Tensor] = []
        self.static_gout: List[torch.Tensor] = []
        self.static_gin: List[Optional[torch.Tensor]] = []
        self.static_pgrads: List[Sequence[torch.Tensor]] = []
        buffers = {k: b.detach().clone() for k, b in module.named_buffers()}
        with torch.cuda.device(device):
            side = torch.cuda.Stream(device)
            side.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(side):
                # eager warm-up: lazy initialisation (kernel attributes, side streams, library handles) must not
                # happen inside a capture
                for _ in range(2):
                    x = sample.detach().clone().requires_grad_(input_requires_grad)
                    out = module(x)
                    torch.autograd.grad(out, ([x] if input_requires_grad else []) + self.params,
                                        torch.ones_like(out), allow_unused=True)
            torch.cuda.current_stream(device).wait_stream(side)
            torch.cuda.synchronize(device)
            pool = torch.cuda.graph_pool_handle()
            # All forward graphs first, every output (and with it the tensors its backward needs) kept alive, then
            # the backward graphs in reverse - the order the pipeline replays them in.  Graphs of a shared pool may
            # reuse memory that an earlier capture has released, so a micro-batch's saved activations must still be
            # referenced while the later micro-batches' forwards are captured, or they would be overwritten by
            # those forwards at replay, before its own backward has run.
            outs = []
            for i in range(n):
                x = sample.detach().clone().requires_grad_(input_requires_grad)
                if i == 0:
                    _mark_shadows_stale(module)
                gf = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gf, pool=pool, stream=side):
                    if self.pull_from is not None:
                        with torch.no_grad():
                            peer_copy(x, self.pull_from[i])
                    out = module(x) if out_slices is None else module(x, out=out_slices[i])
                outs.append(out)
                self.fwd.append(gf)
                self.static_in.append(x)
                self.static_out.append(out.detach())
                self.static_gout.append(torch.empty_like(out))
            self.bwd = [None] * n
            self.static_gin = [None] * n
            self.static_pgrads = [None] * n
            for i in reversed(range(n)):
                gb = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gb, pool=pool, stream=side):
                    grads = torch.autograd.grad(outs[i], ([self.static_in[i]] if input_requires_grad else []) + self.params,
                                                self.static_gout[i], allow_unused=True)
                    if self.push_gin_to is not None and input_requires_grad:
                        with torch.no_grad():
                            peer_copy(self.push_gin_to[i], grads[0].contiguous() if not _dense(grads[0]) else grads[0])
                outs[i] = None
                self.bwd[i] = gb
                self.static_gin[i] = grads[0] if input_requires_grad else None
                self.static_pgrads[i] = list(grads[1:] if input_requires_grad else grads)
            torch.cuda.synchronize(device)
        # warm-up changed the BatchNorm running statistics: put them back
        with torch.no_grad():
            for k, b in module.named_buffers():
                b.copy_(buffers[k])
        self._live = [(p, j) for j, p in enumerate(self.params) if self.static_pgrads[0][j] is not None]
        self._anchor = torch.zeros((), device=device, requires_grad=True)

    def __call__(self, x: torch.Tensor, i: int) -> torch.Tensor:
        # the anchor makes the node part of the autograd graph even when the stage input needs no gradient (the
        # first stage): the parameter gradients are produced inside its backward, not through autograd edges
        return _Replay.apply(x, self._anchor, self, i)

    def accumulate(self, i: int) -> None:
        """``p.grad += `` the static gradients micro-batch ``i``'s backward graph just produced."""
        sg = self.static_pgrads[i]
        with torch.no_grad():
            fresh = [(p, sg[j]) for p, j in self._live if p.grad is None]
            have = [(p, sg[j]) for p, j in self._live if p.grad is not None]
            for p, g in fresh:
                p.grad = g.clone()
            if have:
                torch._foreach_add_([p.grad for p, _ in have], [g for _, g in have])
