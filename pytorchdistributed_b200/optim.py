"""Fused multi-tensor optimizers.

``SGD`` and ``Adam`` are ``torch.optim.Optimizer`` subclasses (same constructor keywords, same
``state_dict`` layout as torch for the fields they share) whose ``step()`` is ONE kernel launch for
all parameters on a CUDA device: a device-resident table of (param, grad, state, bf16-shadow)
pointers plus a block map (``csrc/loss_optim.cu``).  The kernel also refreshes the bf16 shadow copy
that the tensor-core kernels read, so no separate cast pass runs after the update.

Reference call sites: ``torch.optim.SGD(model.parameters(), lr=1e-3)`` at
``R/02 DDP基本概念/ddp_gpus.py:78`` and ``torch.optim.Adam(model.parameters(), lr=1e-3)`` at
``R/03 模型并行/03_model_parallel.ipynb#cell7``.  On CPU tensors the same formulas run in PyTorch.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional

import torch
from torch.optim import Optimizer

from . import _lib


def _shadow_modules(model: Optional[torch.nn.Module]) -> Dict[int, torch.nn.Module]:
    out: Dict[int, torch.nn.Module] = {}
    if model is None:
        return out
    for m in model.modules():
        if hasattr(m, "shadow") and hasattr(m, "mark_shadow_fresh") and not getattr(m, "is_stem", False) \
                and not getattr(m, "small", False):
            out[id(m.weight)] = m
    return out


class _DeviceTable:
    """Pointer table + block map for one device; re-uploaded only when a pointer changes."""

    def __init__(self, device: torch.device) -> None:
        self.device = device
        self.key: Optional[tuple] = None
        self.table: Optional[torch.Tensor] = None
        self.blockmap: Optional[torch.Tensor] = None
        self.n_blocks = 0

    def update(self, rows: List[List[int]]) -> None:
        key = tuple(map(tuple, rows))
        if key == self.key:
            return
        chunk = _lib.call_ret("ptd_opt_chunk_elems")
        bm: List[int] = []
        for ti, r in enumerate(rows):
            for c in range((r[5] + chunk - 1) // chunk):
                bm += [ti, c]
        self.table = torch.tensor(rows, dtype=torch.int64).to(self.device)
        self.blockmap = torch.tensor(bm, dtype=torch.int32).to(self.device)
        self.n_blocks = len(bm) // 2
        self.key = key


class _FusedBase(Optimizer):
    def __init__(self, params, defaults, model: Optional[torch.nn.Module] = None) -> None:
        super().__init__(params, defaults)
        self._shadow_mods = _shadow_modules(model)
        self._tables: Dict[tuple, _DeviceTable] = {}
        # Hyper-parameters the kernels read from DEVICE memory, per (param group, device): [lr] as one float and, for
        # Adam, the step count as one int.  A step captured into a CUDA graph replays kernels, not this Python code:
        # with host scalars the learning rate and Adam's bias corrections would be frozen at their capture-time values.
        self._lr_dev: Dict[tuple, torch.Tensor] = {}
        self._lr_host: Dict[tuple, float] = {}
        self._step_dev: Dict[tuple, torch.Tensor] = {}

    def _lr_tensor(self, gi: int, dev: torch.device, lr: float) -> torch.Tensor:
        key = (gi, dev)
        t = self._lr_dev.get(key)
        if t is None:
            t = self._lr_dev[key] = torch.full((1,), float(lr), device=dev, dtype=torch.float32)
            self._lr_host[key] = float(lr)
        elif self._lr_host[key] != float(lr):
            t.fill_(float(lr))
            self._lr_host[key] = float(lr)
        return t

    def sync_hyper(self) -> None:
        """Push ``param_groups[i]["lr"]`` to the device copies the kernels read.  ``step()`` does this itself; call it
        before replaying a CUDA graph that contains ``step()`` (``Trainer`` does) so an LR schedule takes effect."""
        for (gi, dev), _ in list(self._lr_dev.items()):
            self._lr_tensor(gi, dev, self.param_groups[gi]["lr"])

    def attach_model(self, model: torch.nn.Module) -> "_FusedBase":
        """Let ``step()`` refresh the bf16 shadow weights of ``model`` in the same kernel."""
        self._shadow_mods = _shadow_modules(model)
        return self

    def _rows(self, group_idx: int, params: List[torch.Tensor], n_state: int, state_names: List[str]):
        by_dev: Dict[torch.device, List[List[int]]] = {}
        plist: Dict[torch.device, List[torch.Tensor]] = {}
        for p in params:
            g = p.grad
            if g.dtype != torch.float32 or p.dtype != torch.float32:
                raise RuntimeError("fused optimizers expect fp32 master parameters and gradients")
            if g.stride() != p.stride():
                g = torch.empty_like(p).copy_(g)       # same element order as the parameter
                p.grad = g
            st = self.state[p]
            s = [st[nm].data_ptr() for nm in state_names] + [0] * (2 - n_state)
            mod = self._shadow_mods.get(id(p))
            sh = mod.shadow().data_ptr() if mod is not None else 0
            by_dev.setdefault(p.device, []).append([p.data_ptr(), g.data_ptr(), s[0], s[1], sh, p.numel()])
            plist.setdefault(p.device, []).append(p)
        return by_dev, plist

    def _launch(self, group_idx: int, by_dev, plist, fn) -> None:
        for dev, rows in by_dev.items():
            tab = self._tables.setdefault((group_idx, dev), _DeviceTable(dev))
            tab.update(rows)
            with torch.cuda.device(dev):
                fn(tab, torch.cuda.current_stream(dev).cuda_stream)
            torch.autograd.graph.increment_version(plist[dev])
            for p in plist[dev]:
                mod = self._shadow_mods.get(id(p))
                if mod is not None:
                    mod.mark_shadow_fresh()


class SGD(_FusedBase):
    def __init__(self, params: Iterable, lr: float = 1e-3, momentum: float = 0.0, weight_decay: float = 0.0,
                 model: Optional[torch.nn.Module] = None) -> None:
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay), model)

    @torch.no_grad()
    def step(self, closure=None, grad_scale: float = 1.0):
        loss = closure() if closure is not None else None
        fused = getattr(self, "_ddp_fused", None)
        if fused is not None:
            self._lr_tensor(0, fused.device, self.param_groups[0]["lr"])      # keep the device copy of lr current
            if fused.consume_fused_step():
                # DistributedDataParallel.fuse_optimizer: the gradient all-reduce kernels of the backward that just ran
                # already applied this step (parameters, momentum, bf16 shadows); only the bookkeeping is left
                ps = [p for p in self.param_groups[0]["params"]]
                torch.autograd.graph.increment_version(ps)
                for p in ps:
                    mod = self._shadow_mods.get(id(p))
                    if mod is not None:
                        mod.mark_shadow_fresh()
                return loss
        for gi, group in enumerate(self.param_groups):
            lr, mom, wd = group["lr"], group["momentum"], group["weight_decay"]
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            first = False
            if mom != 0.0:
                for p in ps:
                    st = self.state[p]
                    if "momentum_buffer" not in st:
                        st["momentum_buffer"] = torch.zeros_like(p)
                        first = True
            cuda = [p for p in ps if p.is_cuda]
            cpu = [p for p in ps if not p.is_cuda]
            if cuda:
                by_dev, plist = self._rows(gi, cuda, 1 if mom != 0.0 else 0, ["momentum_buffer"] if mom != 0.0 else [])
                self._launch(gi, by_dev, plist, lambda tab, st: _lib.call(
                    "ptd_multi_sgd_dev", tab.table, tab.blockmap, tab.n_blocks, self._lr_tensor(gi, tab.device, lr),
                    float(mom), float(wd), float(grad_scale), int(first), st))
            for p in cpu:
                g = p.grad * grad_scale
                if wd != 0.0:
                    g = g.add(p, alpha=wd)
                if mom != 0.0:
                    buf = self.state[p]["momentum_buffer"]
                    if first:
                        buf.copy_(g)
                    else:
                        buf.mul_(mom).add_(g)
                    g = buf
                p.add_(g, alpha=-lr)
        return loss


class Adam(_FusedBase):
    def __init__(self, params: Iterable, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, model: Optional[torch.nn.Module] = None) -> None:
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay), model)

    @torch.no_grad()
    def step(self, closure=None, grad_scale: float = 1.0):
        loss = closure() if closure is not None else None
        for gi, group in enumerate(self.param_groups):
            lr, (b1, b2), eps, wd = group["lr"], group["betas"], group["eps"], group["weight_decay"]
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            for p in ps:
                st = self.state[p]
                if "exp_avg" not in st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                st["step"] += 1
            t = self.state[ps[0]]["step"]
            cuda = [p for p in ps if p.is_cuda]
            cpu = [p for p in ps if not p.is_cuda]
            if cuda:
                by_dev, plist = self._rows(gi, cuda, 2, ["exp_avg", "exp_avg_sq"])
                for dev in by_dev:
                    if (gi, dev) not in self._step_dev:      # the kernel increments before it reads: start one behind
                        self._step_dev[(gi, dev)] = torch.full((1,), int(t) - 1, device=dev, dtype=torch.int32)
                self._launch(gi, by_dev, plist, lambda tab, st: _lib.call(
                    "ptd_multi_adam_dev", tab.table, tab.blockmap, tab.n_blocks, self._lr_tensor(gi, tab.device, lr),
                    float(b1), float(b2), float(eps), float(wd), self._step_dev[(gi, tab.device)], float(grad_scale),
                    st))
            for p in cpu:
                st = self.state[p]
                g = p.grad * grad_scale
                if wd != 0.0:
                    g = g.add(p, alpha=wd)
                st["exp_avg"].mul_(b1).add_(g, alpha=1 - b1)
                st["exp_avg_sq"].mul_(b2).addcmul_(g, g, value=1 - b2)
                bias1, bias2 = 1 - b1 ** t, 1 - b2 ** t
                denom = (st["exp_avg_sq"].sqrt() / bias2 ** 0.5).add_(eps)
                p.addcdiv_(st["exp_avg"], denom, value=-lr / bias1)
        return loss

    def step_count(self) -> int:
        """Steps taken so far.  On CUDA the authoritative count lives on the device (it also advances when a captured
        step is replayed, which never runs ``step()``'s Python); reading it synchronises."""
        if self._step_dev:
            return max(int(t.item()) for t in self._step_dev.values())
        steps = [st["step"] for st in self.state.values() if "step" in st]
        return int(max(steps)) if steps else 0

    def state_dict(self):
        if self._step_dev:                       # bring the host mirrors in line with the device count first
            n = self.step_count()
            for st in self.state.values():
                if "step" in st:
                    st["step"] = n
        return super().state_dict()

    def load_state_dict(self, state_dict) -> None:
        super().load_state_dict(state_dict)
        self._step_dev.clear()                   # re-created from the loaded host count at the next step()
