"""Parameter-holding modules whose forward runs on the sm_100a kernels.

Names and shapes of parameters/buffers follow ``torch.nn`` / torchvision exactly, so a
``state_dict`` saved from these modules loads into the reference's models
(``torch.nn.Linear(20, 1)`` in ``R/02 DDP基本概念/ddp_gpus.py:77``, torchvision
``ResNet(Bottleneck,[3,4,6,3])`` in ``R/03 模型并行/03_model_parallel.ipynb#cell4``) and back.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import nn

from ..ops import functional as F_

CL = torch.channels_last


class Conv2d(nn.Module):
    """Bias-free convolution, fp32 master weight (channels_last) + lazily refreshed bf16 shadow."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size: int, stride: int = 1, padding: int = 0,
                 bias: bool = False) -> None:
        super().__init__()
        if bias:
            raise NotImplementedError("ResNet convolutions carry no bias")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.padding = kernel_size, stride, padding
        w = torch.empty(out_channels, in_channels, kernel_size, kernel_size)
        nn.init.kaiming_normal_(w, mode="fan_out", nonlinearity="relu")   # torchvision ResNet init
        self.weight = nn.Parameter(w.contiguous(memory_format=CL))
        self._shadow: Optional[torch.Tensor] = None
        self._shadow_key = None
        self.is_stem = in_channels == 3 and kernel_size == 7 and stride == 2 and padding == 3 and out_channels == 64

    def _apply(self, fn, recurse=True):
        out = super()._apply(fn, recurse)
        self._shadow, self._shadow_key = None, None
        # nn.Module._apply may hand back a differently-strided tensor; the kernels need channels_last.
        if not self.weight.data.is_contiguous(memory_format=CL):
            self.weight.data = self.weight.data.contiguous(memory_format=CL)
        return out

    def mark_shadow_fresh(self) -> None:
        """Called by the fused optimizer after it rewrote weight and shadow in one kernel."""
        self._shadow_key = (self.weight.data_ptr(), self.weight._version)

    def shadow(self) -> torch.Tensor:
        """bf16 copy in the kernel's layout: [Cout][R][S][Cin]; for the stem both packed forms (``F_.stem_pack``)."""
        w = self.weight
        key = (w.data_ptr(), w._version)
        if self._shadow is None or self._shadow_key != key or self._shadow.device != w.device:
            with torch.no_grad():
                if not w.is_cuda:
                    self._shadow = w.detach().to(torch.bfloat16)
                elif self.is_stem:
                    self._shadow = F_.stem_pack(w.detach())
                else:
                    if self._shadow is None or self._shadow.shape != w.shape or self._shadow.device != w.device:
                        self._shadow = torch.empty(w.shape, device=w.device, dtype=torch.bfloat16, memory_format=CL)
                    from .. import _lib
                    with torch.cuda.device(w.device):       # the cast runs on the weight's GPU, whatever is current
                        _lib.call("ptd_cast_f32_to_bf16", w.detach(), self._shadow, w.numel(), F_._stream())
            self._shadow_key = key
        return self._shadow

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.is_stem and x.is_cuda:
            raise RuntimeError("the 7x7 stem is fused with its BatchNorm: call models.resnet.stem_forward / use ResNet.forward")
        return F_.conv2d(x, self.weight, self.shadow(), self.stride, self.padding)

    def extra_repr(self) -> str:
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, stride={self.stride}, "
                f"padding={self.padding}, bias=False")


class BatchNorm2d(nn.Module):
    def __init__(self, num_features: int, eps: float = 1e-5, momentum: float = 0.1) -> None:
        super().__init__()
        self.num_features, self.eps, self.momentum = num_features, eps, momentum
        self.weight = nn.Parameter(torch.ones(num_features))
        self.bias = nn.Parameter(torch.zeros(num_features))
        self.register_buffer("running_mean", torch.zeros(num_features))
        self.register_buffer("running_var", torch.ones(num_features))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))
        self.fuse_relu = False      # set by a container that wants BN+ReLU in one pass

    def eval_affine(self):
        scale = self.weight * torch.rsqrt(self.running_var + self.eps)
        shift = self.bias - self.running_mean * scale
        return scale.float().contiguous(), shift.float().contiguous()

    def forward(self, y: torch.Tensor, residual: Optional[torch.Tensor] = None, relu: Optional[bool] = None):
        relu = self.fuse_relu if relu is None else relu
        if self.training:
            self.num_batches_tracked += 1
            return F_.batch_norm_act(y, self.weight, self.bias, self.running_mean, self.running_var, residual,
                                     self.eps, self.momentum, relu)
        scale, shift = self.eval_affine()
        return F_.affine_act(y, scale, shift, residual, relu)

    def extra_repr(self) -> str:
        return f"{self.num_features}, eps={self.eps}, momentum={self.momentum}"


def conv_bn_act(x, conv: Conv2d, bn: BatchNorm2d, relu: bool = True, residual: Optional[torch.Tensor] = None):
    """conv -> BN (batch stats from the conv epilogue) -> (+residual) -> (ReLU): two kernels forward."""
    if bn.training:
        with torch.no_grad():
            bn.num_batches_tracked += 1
        return F_.conv_bn_act(x, conv.weight, conv.shadow(), bn.weight, bn.bias, bn.running_mean, bn.running_var,
                              residual, conv.stride, conv.padding, bn.eps, bn.momentum, relu,
                              stem=conv.is_stem and x.is_cuda)
    if conv.is_stem and x.is_cuda:
        y, _ = F_.stem_fprop_raw(x, conv.shadow())
    else:
        y = F_.conv2d(x, conv.weight, conv.shadow(), conv.stride, conv.padding)
    scale, shift = bn.eval_affine()
    return F_.affine_act(y, scale, shift, residual, relu)


class MaxPool2d(nn.Module):
    """3x3 / stride 2 / pad 1 (the only pooling window ResNet uses)."""

    def __init__(self, kernel_size: int = 3, stride: int = 2, padding: int = 1) -> None:
        super().__init__()
        assert (kernel_size, stride, padding) == (3, 2, 1)

    def forward(self, x):
        return F_.max_pool_3x3_s2(x)


class GlobalAvgPool(nn.Module):
    """AdaptiveAvgPool2d((1,1)); returns [N, C, 1, 1] like torch so ``x.view(x.size(0), -1)`` works."""

    def forward(self, x):
        return F_.global_avg_pool(x).view(x.size(0), x.size(1), 1, 1)


class Linear(nn.Module):
    """Dense layer.  Large shapes go to the tcgen05 GEMM, tutorial-sized ones (<= 64 wide) to the
    fp32 CUDA-core kernel."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True) -> None:
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.empty(out_features)) if bias else None
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if bias:
            bound = 1 / math.sqrt(in_features)
            nn.init.uniform_(self.bias, -bound, bound)
        self._shadow, self._shadow_key = None, None
        self.small = in_features <= 64 and out_features <= 64

    def _apply(self, fn, recurse=True):
        out = super()._apply(fn, recurse)
        self._shadow, self._shadow_key = None, None
        return out

    def mark_shadow_fresh(self) -> None:
        self._shadow_key = (self.weight.data_ptr(), self.weight._version)

    def shadow(self) -> torch.Tensor:
        w = self.weight
        key = (w.data_ptr(), w._version)
        if self._shadow is None or self._shadow_key != key:
            with torch.no_grad():
                if w.is_cuda:
                    if self._shadow is None or self._shadow.device != w.device:
                        self._shadow = torch.empty(w.shape, device=w.device, dtype=torch.bfloat16)
                    from .. import _lib
                    with torch.cuda.device(w.device):
                        _lib.call("ptd_cast_f32_to_bf16", w.detach().contiguous(), self._shadow, w.numel(), F_._stream())
                else:
                    self._shadow = w.detach().to(torch.bfloat16)
            self._shadow_key = key
        return self._shadow

    def forward(self, x: torch.Tensor, relu: bool = False, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``out``: preallocated fp32 ``[rows, out_features]`` block to write the result into (tensor-core path only)."""
        if self.small:
            return F_.small_linear(x, self.weight, self.bias, relu)
        res = F_.linear(x.reshape(x.size(0), -1), self.weight, self.shadow(), self.bias, out=None if relu else out)
        return torch.relu(res) if relu else res

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}"
