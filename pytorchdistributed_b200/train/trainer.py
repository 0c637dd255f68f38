"""``Trainer`` - the training loop of the reference's DDP scripts, plus rank-0 checkpoints.

Same constructor surface and per-epoch behaviour as ``R/02 DDP基本概念/ddp_gpus.py:25-55`` (spawn:
``Trainer(model, train_dataloader, optimizer, gpu_id)``) and ``ddp_gpus_torchrun.py:21-54``
(torchrun: ``gpu_id`` read from ``LOCAL_RANK``):

* the model is moved to the device and wrapped in ``DistributedDataParallel``;
* ``_run_epoch`` peeks one batch to print ``[GPU: r] Epoch: e | Batchsize: b | Steps: n`` (exact
  format, golden-tested), calls ``sampler.set_epoch(epoch)``, then ``.to(device)`` + ``_run_batch``;
* ``_run_batch`` = ``zero_grad -> forward -> cross_entropy -> backward -> step``.

Additions (defaults keep the reference behaviour): ``save_every``/``snapshot_path`` write
``model.module.state_dict()`` from rank 0 (keys without the ``module.`` prefix, so the file loads
into the reference's own model definitions); ``resume`` reloads it; ``loss_fn`` swaps the loss
(the reference's float-target cross-entropy on one logit is identically zero - SURVEY.md 2.6 quirk 1).
"""
from __future__ import annotations

import os
from typing import Callable, Optional

import torch
import torch.distributed as dist
from torch.utils.data import DataLoader

from ..ops import functional as K
from ..parallel.ddp import DistributedDataParallel as DDP
from ..utils.metrics import JsonlWriter, StepMeter
from ..utils.nvtx import nvtx_range


class Trainer:
    def __init__(self, model: torch.nn.Module, train_dataloader: DataLoader, optimizer: torch.optim.Optimizer,
                 gpu_id: Optional[int] = None, *, save_every: int = 0, snapshot_path: str = "snapshot.pt",
                 loss_fn: Optional[Callable] = None, resume: bool = False, ddp_kwargs: Optional[dict] = None,
                 quiet: bool = False, cuda_graph: bool = False, graph_warmup: int = 3,
                 metrics_path: Optional[str] = None, fuse_optimizer: bool = False) -> None:
        self.gpu_id = int(os.environ.get("LOCAL_RANK", 0)) if gpu_id is None else gpu_id
        self.device = torch.device("cuda", self.gpu_id) if torch.cuda.is_available() else torch.device("cpu")
        self.model = model.to(self.device)
        self.train_dataloader = train_dataloader
        self.optimizer = optimizer
        self.save_every = save_every
        self.snapshot_path = snapshot_path
        self.loss_fn = loss_fn or K.cross_entropy
        self.quiet = quiet
        self.epochs_run = 0
        self.last_loss: Optional[torch.Tensor] = None
        self.meter = StepMeter(self.device)
        self._metrics = JsonlWriter(metrics_path if self._is_rank0() else None)
        if resume and os.path.exists(snapshot_path):
            self._load_snapshot(snapshot_path)
        kw = dict(ddp_kwargs or {})
        self.model = DDP(self.model, device_ids=[self.gpu_id] if self.device.type == "cuda" else None, **kw)
        if hasattr(optimizer, "attach_model"):
            optimizer.attach_model(self.model.module)
        # SGD step applied by the gradient all-reduce kernels themselves (DistributedDataParallel.fuse_optimizer);
        # silently stays unfused where that is not possible (one rank, NCCL / gloo backend, another optimizer)
        self.fused_optimizer = bool(fuse_optimizer) and self.model.fuse_optimizer(optimizer)
        # CUDA-graph replay of the whole step (zero_grad .. optimizer.step) once shapes have settled
        self.cuda_graph = bool(cuda_graph) and self.device.type == "cuda"
        self._graph_warmup = max(int(graph_warmup), 2)
        self._eager_steps = 0
        self._graph = None
        self._graph_failed = False

    # -- one optimisation step ----------------------------------------------------------------
    def _step_eager(self, xs, ys):
        self.optimizer.zero_grad()
        output = self.model(xs)
        loss = self.loss_fn(output, ys)
        loss.backward()
        self.optimizer.step()
        return loss.detach()

    def _run_batch(self, xs, ys):
        if self.cuda_graph and not self._graph_failed:
            g = self._graph
            if g is not None and g["x"].shape == xs.shape and g["y"].shape == ys.shape and g["x"].dtype == xs.dtype:
                g["x"].copy_(xs, non_blocking=True)
                g["y"].copy_(ys, non_blocking=True)
                if hasattr(self.optimizer, "sync_hyper"):
                    self.optimizer.sync_hyper()          # an LR schedule reaches the replayed kernels through device memory
                g["graph"].replay()
                self.last_loss = g["loss"]
                return self.last_loss
            if g is None and self._eager_steps >= self._graph_warmup:
                self._capture(xs, ys)
                if self._graph is not None:
                    return self._run_batch(xs, ys)
        self.last_loss = self._step_eager(xs, ys)
        self._eager_steps += 1
        return self.last_loss

    def _capture(self, xs, ys) -> None:
        """Record one full training step (forward, loss, backward with the overlapped gradient all-reduce,
        optimizer) into a CUDA graph on static input buffers; later steps are a copy + one graph launch."""
        try:
            sx, sy = xs.clone(), ys.clone()
            torch.cuda.synchronize(self.device)
            from .. import _lib
            graph = torch.cuda.CUDAGraph()
            l0 = _lib.LAUNCHES
            with torch.cuda.graph(graph):
                loss = self._step_eager(sx, sy)
            torch.cuda.synchronize(self.device)
            # "launches": native entry-point calls recorded into the graph (each replay re-issues them on the device)
            self._graph = {"graph": graph, "x": sx, "y": sy, "loss": loss, "launches": _lib.LAUNCHES - l0}
        except Exception as e:  # noqa: BLE001 - fall back to eager, keep training
            self._graph_failed = True
            self._graph = None
            torch.cuda.synchronize(self.device)
            # always said, also when `quiet`: a silent fall-back to eager would hide a 2x slower training loop
            print(f"[GPU: {self.gpu_id}] CUDA-graph capture failed, staying eager: {type(e).__name__}: {e}", flush=True)

    def _run_epoch(self, epoch: int):
        batch_size = len(next(iter(self.train_dataloader))[0])
        if not self.quiet:
            print(f"[GPU: {self.gpu_id}] Epoch: {epoch} | Batchsize: {batch_size} | Steps: {len(self.train_dataloader)}",
                  flush=True)
        sampler = getattr(self.train_dataloader, "sampler", None)
        if hasattr(sampler, "set_epoch"):
            sampler.set_epoch(epoch)
        for step, (xs, ys) in enumerate(self.train_dataloader):
            xs = xs.to(self.device, non_blocking=True)
            ys = ys.to(self.device, non_blocking=True)
            with nvtx_range(f"train_step/e{epoch}"):
                self._run_batch(xs, ys)
            stats = self.meter.step(len(xs))
            if stats is not None:
                self._metrics.write(epoch=epoch, step=step, loss=float(self.last_loss), **stats)

    def run_batches(self, batches):
        """Train on an iterable of HOST batches ``(xs, ys)`` (pinned memory recommended) and return the per-step
        losses as Python floats.  The copy of batch i+1 runs on a side stream while step i computes, and the loss of
        step i is read back (device -> host) while step i+1 runs, so neither transfer stalls the GPU."""
        if self.device.type != "cuda":
            return [float(self._run_batch(xs, ys)) for xs, ys in batches]
        if not hasattr(self, "_copy_stream"):
            self._copy_stream = torch.cuda.Stream(self.device)
        it = iter(batches)
        cur_stream = torch.cuda.current_stream(self.device)

        def stage(batch):
            with torch.cuda.stream(self._copy_stream):
                xs = batch[0].to(self.device, non_blocking=True)
                ys = batch[1].to(self.device, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self._copy_stream)
            return xs, ys, ev

        losses, pending = [], None
        nxt = next(it, None)
        staged = stage(nxt) if nxt is not None else None
        while staged is not None:
            xs, ys, ev = staged
            cur_stream.wait_event(ev)
            xs.record_stream(cur_stream)
            ys.record_stream(cur_stream)
            loss = self._run_batch(xs, ys)
            nxt = next(it, None)
            staged = stage(nxt) if nxt is not None else None       # overlaps with the step just launched
            host = torch.empty(1, dtype=torch.float32, pin_memory=True)
            host.copy_(loss.detach().reshape(1).float(), non_blocking=True)
            done = torch.cuda.Event()
            done.record(cur_stream)
            if pending is not None:                                  # read the previous step's loss now
                pending[1].synchronize()
                losses.append(float(pending[0]))
            pending = (host, done)
        if pending is not None:
            pending[1].synchronize()
            losses.append(float(pending[0]))
        return losses

    def train(self, max_epoch: int):
        for epoch in range(self.epochs_run, max_epoch):
            self._run_epoch(epoch)
            if self.save_every and (epoch + 1) % self.save_every == 0:
                self._save_snapshot(epoch)

    # -- checkpoints ----------------------------------------------------------------------------
    def _is_rank0(self) -> bool:
        return (not dist.is_initialized()) or dist.get_rank() == 0

    def _save_snapshot(self, epoch: int) -> None:
        if self._is_rank0():
            inner = self.model.module if hasattr(self.model, "module") else self.model
            state = {k: v.detach().cpu().contiguous() for k, v in inner.state_dict().items()}
            torch.save({"MODEL_STATE": state, "EPOCHS_RUN": epoch + 1}, self.snapshot_path)
            if not self.quiet:
                print(f"Epoch {epoch} | Training snapshot saved at {self.snapshot_path}", flush=True)
        if dist.is_initialized():
            dist.barrier()

    def _load_snapshot(self, path: str) -> None:
        snap = torch.load(path, map_location=self.device)
        self.model.load_state_dict(snap["MODEL_STATE"])
        self.epochs_run = int(snap.get("EPOCHS_RUN", 0))
        if not self.quiet:
            print(f"[GPU: {self.gpu_id}] Resuming training from snapshot at epoch {self.epochs_run}", flush=True)


def save_model_state(model: torch.nn.Module, path: str) -> None:
    """Plain ``state_dict`` file (no wrapper dict): ``torch.nn.Linear(20,1).load_state_dict(torch.load(path))``."""
    inner = model.module if hasattr(model, "module") else model
    torch.save({k: v.detach().cpu().contiguous() for k, v in inner.state_dict().items()}, path)
