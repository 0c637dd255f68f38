"""Process-per-stage pipeline parallelism with GPipe and 1F1B schedules (+ optional activation recompute).

The reference only *describes* these schedules (``R/03 模型并行/03_model_parallel.ipynb#cell14-15``: GPipe
re-materialisation, 1F1B / PipeDream-flush) and implements a single-process two-stage forward overlap.  This module
is the one-process-per-GPU form: every rank owns one stage, activations travel forward and activation gradients
backward with ``torch.distributed`` point-to-point (NCCL send/recv on a GPU box - stream-ordered, overlapping the
next micro-batch's compute; gloo on a CPU box, which is what ``tests/test_cpu_pipeline_mp.py`` uses).

* ``schedule_ops(schedule, stage, n_stages, n_micro)`` - the per-stage order of ``("F", i)`` / ``("B", i)`` steps.
  ``"gpipe"``: all forwards, then all backwards.  ``"1f1b"``: ``n_stages - stage - 1`` warm-up forwards, then one
  forward / one backward alternately, then the remaining backwards - at most ``n_stages - stage`` activations
  alive instead of ``n_micro``.
* ``PipelineStage.train_step`` runs one optimisation step's forward/backward over ``n_micro`` micro-batches
  (``split_size`` rows each) and leaves summed gradients in ``.grad``; the caller steps its optimizer.
  ``recompute=True`` keeps only each micro-batch's stage input and re-runs the stage forward right before its
  backward (GPipe-style re-materialisation).
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
from torch import nn

Op = Tuple[str, int]


def schedule_ops(schedule: str, stage: int, n_stages: int, n_micro: int) -> List[Op]:
    if schedule == "gpipe":
        return [("F", i) for i in range(n_micro)] + [("B", i) for i in range(n_micro)]
    if schedule != "1f1b":
        raise ValueError(f"unknown schedule {schedule!r}")
    warm = min(n_stages - stage - 1, n_micro)
    ops: List[Op] = [("F", i) for i in range(warm)]
    f, b = warm, 0
    while f < n_micro:
        ops.append(("F", f)); f += 1
        ops.append(("B", b)); b += 1
    while b < n_micro:
        ops.append(("B", b)); b += 1
    return ops


def max_live_activations(ops: Sequence[Op]) -> int:
    live = peak = 0
    for kind, _ in ops:
        live += 1 if kind == "F" else -1
        peak = max(peak, live)
    return peak


class PipelineStage:
    def __init__(self, module: nn.Module, stage: int, n_stages: int, device: Optional[torch.device] = None,
                 group=None, loss_fn: Optional[Callable] = None, schedule: str = "1f1b", recompute: bool = False) -> None:
        self.module, self.stage, self.n_stages = module, stage, n_stages
        self.group = group
        self.device = device or next(module.parameters()).device
        self.loss_fn = loss_fn
        self.schedule = schedule
        self.recompute = recompute
        self.first, self.last = stage == 0, stage == n_stages - 1
        self._act_meta = None           # (shape tail, dtype) of the incoming activation, learnt on the first step
        self._pending: List = []        # outstanding isend handles (buffers must outlive them)

    # -- point-to-point helpers -------------------------------------------------------------------------
    def _peer(self, delta: int) -> int:
        r = self.stage + delta
        return dist.get_global_rank(self.group, r) if self.group is not None else r

    def _send(self, t: torch.Tensor, delta: int) -> None:
        t = t.detach().contiguous()
        self._pending.append((dist.isend(t, self._peer(delta), group=self.group), t))

    def _recv(self, shape, dtype, delta: int) -> torch.Tensor:
        buf = torch.empty(shape, dtype=dtype, device=self.device)
        dist.recv(buf, self._peer(delta), group=self.group)
        return buf

    def _handshake(self, out: Optional[torch.Tensor], rows: int) -> None:
        """First step only: tell the next stage what a micro-batch activation looks like."""
        if not self.last and out is not None:
            meta = torch.tensor([out.dim()] + list(out.shape[1:]) + [0] * (7 - out.dim()) +
                                [{torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}[out.dtype]],
                                dtype=torch.int64, device=self.device)
            dist.send(meta, self._peer(+1), group=self.group)

    def _learn_meta(self) -> None:
        meta = torch.empty(8, dtype=torch.int64, device=self.device)
        dist.recv(meta, self._peer(-1), group=self.group)
        m = meta.tolist()
        self._act_meta = (tuple(m[1:m[0]]), {0: torch.float32, 1: torch.bfloat16, 2: torch.float16}[m[7]])

    # -- one optimisation step -----------------------------------------------------------------------------
    def train_step(self, inputs: Optional[torch.Tensor], targets: Optional[torch.Tensor], split_size: int,
                   batch_rows: Optional[int] = None):
        """``inputs`` is needed on the first stage, ``targets`` on the last, ``batch_rows`` on every other stage
        (so each stage can derive the micro-batch sizes).  Returns the mean loss on the last stage, else ``None``."""
        rows_total = inputs.size(0) if inputs is not None else (targets.size(0) if targets is not None else batch_rows)
        sizes = [min(split_size, rows_total - s) for s in range(0, rows_total, split_size)]
        n_micro = len(sizes)
        starts = [sum(sizes[:i]) for i in range(n_micro)]
        ops = schedule_ops(self.schedule, self.stage, self.n_stages, n_micro)
        saved_in, saved_out = {}, {}
        loss_sum = torch.zeros((), device=self.device)
        for kind, i in ops:
            rows = sizes[i]
            if kind == "F":
                if self.first:
                    x = inputs[starts[i]:starts[i] + rows].to(self.device)
                else:
                    if self._act_meta is None:
                        self._learn_meta()
                    tail, dtype = self._act_meta
                    x = self._recv((rows,) + tail, dtype, -1).requires_grad_()
                if self.recompute:
                    with torch.no_grad():
                        out = self.module(x)
                else:
                    out = self.module(x)
                if not self.last and not getattr(self, "_meta_sent", False):
                    self._handshake(out, rows)
                    self._meta_sent = True
                saved_in[i] = x
                if self.last:
                    y = targets[starts[i]:starts[i] + rows].to(self.device)
                    if self.recompute:
                        saved_out[i] = y
                    else:
                        saved_out[i] = self.loss_fn(out, y) * (rows / rows_total)
                        loss_sum += saved_out[i].detach()
                else:
                    self._send(out, +1)
                    saved_out[i] = None if self.recompute else out
            else:
                x = saved_in.pop(i)
                if self.recompute:
                    out = self.module(x)                     # re-materialise the stage's activations
                    if self.last:
                        loss = self.loss_fn(out, saved_out.pop(i)) * (rows / rows_total)
                        loss_sum += loss.detach()
                        loss.backward()
                    else:
                        saved_out.pop(i, None)
                        g = self._recv(out.shape, out.dtype, +1)
                        out.backward(g)
                else:
                    out = saved_out.pop(i)
                    if self.last:
                        out.backward()
                    else:
                        g = self._recv(out.shape, out.dtype, +1)
                        out.backward(g)
                if not self.first:
                    self._send(x.grad, -1)
        for h, _ in self._pending:
            h.wait()
        self._pending.clear()
        return loss_sum if self.last else None
