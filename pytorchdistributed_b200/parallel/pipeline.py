"""Model-parallel and micro-batch pipeline-parallel ResNet-50 over two GPUs of one process.

* ``ModelParallelResNet50``    - the reference's layer split (``R/03 模型并行/03_model_parallel.ipynb#cell5``):
  ``seq1`` = conv1..layer2 on the first device, ``seq2`` = layer3..avgpool and ``fc`` on the second;
  ``forward`` = ``seq1(x) -> transfer -> seq2 -> fc``.
* ``PipelineParallelResNet50(split_size=20)`` (``#cell12``): the batch is cut into ``split_size`` micro-batches
  and the two stages overlap - stage 2 works on micro-batch i while stage 1 computes i+1.  Uneven last
  split allowed (``torch.split`` semantics).  Backward is plain autograd over the whole graph (all forward,
  then all backward, GPipe-like without recompute), as in the reference.

What changes underneath: activations and activation gradients cross the stage boundary as bf16 NHWC with an
NCCL send/recv pair on per-device side streams (``comm/nccl_p2p.py``), ordered against the compute streams
with events, so the copy of micro-batch i overlaps the compute of i+1 on both devices; the stages run on
the sm_100a kernels.  ``transfer="p2p"`` swaps NCCL for our device-issued peer copy kernel.
Unlike the reference class this one does not double-register sub-modules (quirk 5 in SURVEY.md 2.6):
``state_dict()`` has torchvision's 320 keys; ``reference_state_dict()`` adds the ``seq1.*``/``seq2.*`` aliases.
"""
from __future__ import annotations

from typing import List, Optional

import torch
from torch import nn

from .. import _lib
from ..comm import p2p as _p2p  # noqa: F401
from ..models.resnet import Bottleneck, ResNet
from ..ops import functional as F_

_GROUPS = {}


def _group(d0: int, d1: int):
    from ..comm.nccl_p2p import LocalNcclGroup
    key = (d0, d1)
    if key not in _GROUPS:
        _GROUPS[key] = LocalNcclGroup([d0, d1])
    return _GROUPS[key]


class _Transfer(torch.autograd.Function):
    """Move a tensor to another device on side streams; the gradient travels back the same way."""

    @staticmethod
    def forward(ctx, x: torch.Tensor, dst_index: int, mode: str):
        ctx.src_index, ctx.dst_index, ctx.mode = x.device.index, dst_index, mode
        return _Transfer._move(x, dst_index, mode)

    @staticmethod
    def backward(ctx, g: torch.Tensor):
        return _Transfer._move(g, ctx.src_index, ctx.mode), None, None

    @staticmethod
    def _move(t: torch.Tensor, dst_index: int, mode: str) -> torch.Tensor:
        src_dev, dst_dev = t.device, torch.device("cuda", dst_index)
        cl = t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last) and not t.is_contiguous()
        src = t if (t.is_contiguous() or cl) else t.contiguous()
        with torch.cuda.device(dst_dev):
            out = torch.empty_like(src, device=dst_dev)         # same strides (keeps NHWC)
        if mode == "nccl":
            done = _group(min(src_dev.index, dst_index), max(src_dev.index, dst_index)).send_recv(
                _flat(src), _flat(out))
            torch.cuda.current_stream(dst_dev).wait_event(done)
        else:
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(src_dev))
            with torch.cuda.device(dst_dev):
                st = torch.cuda.current_stream(dst_dev)
                st.wait_event(ready)
                nb = src.numel() * src.element_size()
                if nb % 16 == 0:
                    _lib.call("ptd_peer_copy", out, src, nb, st.cuda_stream)   # the consumer pulls over NVLink
                else:
                    out.copy_(src, non_blocking=True)
            src.record_stream(torch.cuda.current_stream(dst_dev))
        return out


def _flat(t: torch.Tensor) -> torch.Tensor:
    return torch.as_strided(t, (t.numel(),), (1,), t.storage_offset())


class ModelParallelResNet50(ResNet):
    def __init__(self, num_classes: int = 1000, devices=("cuda:0", "cuda:1"), transfer: str = "nccl") -> None:
        super().__init__(block=Bottleneck, layers=[3, 4, 6, 3], num_classes=num_classes)
        self.devices = [torch.device(d) for d in devices]
        self.transfer = transfer
        self._on_gpu = torch.cuda.is_available() and all(d.type == "cuda" for d in self.devices)
        seq1, seq2, fc = self.stage_modules()
        # plain attributes (not registered a second time): state_dict keeps torchvision's keys
        object.__setattr__(self, "seq1", seq1)
        object.__setattr__(self, "seq2", seq2)
        if self._on_gpu:
            seq1.to(self.devices[0])
            seq2.to(self.devices[1])
            fc.to(self.devices[1])
            d0, d1 = self.devices[0].index, self.devices[1].index
            if d0 != d1 and _lib.available():
                # our copy kernels address the other stage's memory directly: both directions must be mapped (torch
                # only enables peer access lazily for its own copies)
                _lib.call("ptd_p2p_enable_peer_access", d0, d1)
                _lib.call("ptd_p2p_enable_peer_access", d1, d0)

    def _boundary(self, x: torch.Tensor) -> torch.Tensor:
        if not self._on_gpu:
            return x
        return _Transfer.apply(x, self.devices[1].index, self.transfer)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self._boundary(self.seq1(x))
        x = self.seq2(x)
        return self.fc(x.view(x.size(0), -1))

    def reference_state_dict(self):
        """state_dict with the alias keys the reference class produces by registering modules twice."""
        sd = dict(self.state_dict())
        for k, v in list(sd.items()):
            head = k.split(".")[0]
            m1 = {"conv1": "seq1.0", "bn1": "seq1.1", "layer1": "seq1.4", "layer2": "seq1.5"}
            m2 = {"layer3": "seq2.0", "layer4": "seq2.1"}
            if head in m1:
                sd[m1[head] + k[len(head):]] = v
            elif head in m2:
                sd[m2[head] + k[len(head):]] = v
        return sd


class _Stage2(nn.Module):
    """layer3..avgpool + fc as one module (the second pipeline stage)."""

    def __init__(self, seq2: nn.Module, fc: nn.Module) -> None:
        super().__init__()
        self.seq2, self.fc = seq2, fc

    def forward(self, x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        x = self.seq2(x)
        return _fc_into(self.fc, x.view(x.size(0), -1), out)


def _fc_into(fc: nn.Module, feat: torch.Tensor, out: Optional[torch.Tensor]) -> torch.Tensor:
    """Classifier of one micro-batch, written straight into its rows of the pipeline's output tensor when the layer
    can do that (our tensor-core ``Linear`` on CUDA); any other layer returns a fresh tensor and is copied in later."""
    if out is not None and feat.is_cuda and getattr(fc, "small", True) is False:
        return fc(feat, out=out)
    return fc(feat)


class PipelineParallelResNet50(ModelParallelResNet50):
    """``cuda_graph=True``: in training mode every full-size micro-batch of either stage replays a captured
    (forward, backward) graph pair (``parallel/graphed.py``) instead of issuing its kernels from Python; graphs are
    built on the first call for that batch size, an uneven last split runs eagerly."""

    def __init__(self, split_size: int = 20, *args, cuda_graph: bool = False, **kwargs) -> None:
        super().__init__(*args, **kwargs)
        self.split_size = split_size
        self.cuda_graph = cuda_graph and self._on_gpu
        self._graphs = None          # (n_full, stage-1 GraphedStage, stage-2 GraphedStage)

    def _graphed(self, x: torch.Tensor):
        from .graphed import GraphedStage
        n_full = x.size(0) // self.split_size
        if self._graphs is None or self._graphs[0] != n_full or self._graphs[1].static_in[0].shape[1:] != x.shape[1:]:
            if n_full == 0:
                return None
            sample = x[:self.split_size].detach().to(self.devices[0])
            g1 = GraphedStage(self.seq1, self.devices[0], sample, n_full, input_requires_grad=False)
            o = g1.static_out[0]                 # shape / strides of the activation that crosses the boundary
            act = torch.empty_strided(o.shape, o.stride(), dtype=o.dtype, device=self.devices[1]).normal_()
            stage2 = _Stage2(self.seq2, self.fc)
            # every micro-batch's logits land in their rows of ONE output tensor (room for an eager tail split)
            with torch.cuda.device(self.devices[1]):
                self._out_buf = torch.empty(((n_full + 1) * self.split_size, self.fc.out_features),
                                            dtype=torch.float32, device=self.devices[1])
            slices = [self._out_buf[i * self.split_size:(i + 1) * self.split_size] for i in range(n_full)]
            if self.transfer == "p2p":
                # the stage boundary rides inside the replayed work: stage 2's forward graph pulls the activation out
                # of stage 1's static output, its backward graph pushes the activation gradient into stage 1's buffer
                g2 = GraphedStage(stage2, self.devices[1], act, n_full, input_requires_grad=True,
                                  pull_from=g1.static_out, push_gin_to=g1.static_gout, out_slices=slices)
            else:
                g2 = GraphedStage(stage2, self.devices[1], act, n_full, input_requires_grad=True, out_slices=slices)
            self._graphs = (n_full, g1, g2)
        return self._graphs

    def _forward_graphed(self, x: torch.Tensor) -> torch.Tensor:
        n_full, g1, g2 = self._graphs
        splits = list(x.split(self.split_size, dim=0))

        def stage1(i):
            s = splits[i]
            return g1(s, i) if i < n_full else self.seq1(s)

        fused_boundary = g2.pull_from is not None
        rows = [s.size(0) for s in splits]
        starts = [sum(rows[:i]) for i in range(len(rows))]
        out_buf = self._out_buf[:x.size(0)]

        def stage2(a, i):
            if i < n_full:
                return g2(a, i)
            a = self.seq2(a)
            return _fc_into(self.fc, a.view(a.size(0), -1), out_buf[starts[i]:starts[i] + rows[i]])

        def boundary(a, i):
            # graph-replayed micro-batches carry the transfer inside stage 2's graphs
            return a if (fused_boundary and i < n_full) else self._boundary(a)

        ret: List[torch.Tensor] = []
        prev = boundary(stage1(0), 0)
        for i in range(1, len(splits)):
            ret.append(stage2(prev, i - 1))
            prev = boundary(stage1(i), i)
        ret.append(stage2(prev, len(splits) - 1))
        # (the returned tensor aliases the stage's static output buffer: the next forward overwrites it)
        return F_.gather_rows(out_buf, ret)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.cuda_graph and self.training and torch.is_grad_enabled() and self._graphed(x) is not None:
            return self._forward_graphed(x)
        # Every micro-batch's classifier writes its [b, classes] rows of ONE preallocated output (no concatenation pass
        # at the end, unlike the reference's torch.cat, notebook 03 cell 12).
        parts = list(x.split(self.split_size, dim=0))
        out_dev = self.devices[1] if self._on_gpu else x.device
        out = torch.empty((x.size(0), self.fc.out_features), dtype=torch.float32, device=out_dev)
        ret: List[torch.Tensor] = []
        row = 0

        def head(feat: torch.Tensor) -> None:
            nonlocal row
            ret.append(_fc_into(self.fc, feat.view(feat.size(0), -1), out[row:row + feat.size(0)]))
            row += feat.size(0)

        s_prev = self._boundary(self.seq1(parts[0]))
        for s_next in parts[1:]:
            head(self.seq2(s_prev))                        # A. the previous micro-batch runs on the second device
            s_prev = self._boundary(self.seq1(s_next))     # B. the next one on the first, concurrently; its transfer
        head(self.seq2(s_prev))                            #    rides the side streams
        return F_.gather_rows(out, ret)
