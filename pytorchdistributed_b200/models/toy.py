"""Tutorial-sized models.

* ``ToyLinear``  - the ``torch.nn.Linear(20, 1)`` of ``R/02 DDP基本概念/ddp_gpus.py:77``
  (state_dict keys ``weight``, ``bias``).
* ``MLP``        - the 4-layer ``Model`` of ``R/01 数据并行/01_multi_gpus_data_parallelism.ipynb#cell4``:
  Linear(10,20)-ReLU-Linear(20,20)-ReLU-Linear(20,20)-ReLU-Linear(20,5) under the attribute ``net``
  (keys ``net.0.weight`` ... ``net.6.bias``).
"""
from __future__ import annotations

import ctypes

import torch
from torch import nn

from .. import _lib
from .layers import Linear


class ToyLinear(Linear):
    def __init__(self, in_features: int = 20, out_features: int = 1) -> None:
        super().__init__(in_features, out_features, bias=True)


class MLP(nn.Module):
    def __init__(self, input_size: int = 10, hidden_size: int = 20, output_size: int = 5) -> None:
        super().__init__()
        self.net = nn.Sequential(
            Linear(input_size, hidden_size), nn.ReLU(),
            Linear(hidden_size, hidden_size), nn.ReLU(),
            Linear(hidden_size, hidden_size), nn.ReLU(),
            Linear(hidden_size, output_size),
        )

    def _linears(self):
        return [m for m in self.net if isinstance(m, Linear)]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        lins = self._linears()
        if x.is_cuda and not torch.is_grad_enabled():
            return self.fused_forward(x)
        for i, lin in enumerate(lins):
            x = lin(x, relu=i + 1 < len(lins))      # ReLU fused into the layer kernel
        return x

    @torch.no_grad()
    def fused_forward(self, x: torch.Tensor) -> torch.Tensor:
        """All layers in one kernel (weights in shared memory) - the notebook's forward-only loop."""
        lins = self._linears()
        x = x.float().contiguous()
        dims = [lins[0].in_features] + [l.out_features for l in lins]
        y = torch.empty((x.size(0), dims[-1]), device=x.device, dtype=torch.float32)
        wp = (ctypes.c_void_p * len(lins))(*[l.weight.data_ptr() for l in lins])
        bp = (ctypes.c_void_p * len(lins))(*[l.bias.data_ptr() for l in lins])
        dm = (ctypes.c_int * len(dims))(*dims)
        _lib.call("ptd_mlp_fwd", x, y, ctypes.cast(wp, ctypes.c_void_p), ctypes.cast(bp, ctypes.c_void_p),
                  ctypes.cast(dm, ctypes.c_void_p), len(lins), x.size(0), torch.cuda.current_stream().cuda_stream)
        return y
