"""Autograd-level ops on top of the sm_100a kernel library.

Activations are ``torch.bfloat16`` tensors with logical shape ``[N, C, H, W]`` in
``torch.channels_last`` memory format, i.e. NHWC in memory - the layout the tensor-core kernels
read with TMA / 128-byte row gathers.  Convolution weights are fp32 masters with logical shape
``[Cout, Cin, R, S]`` in channels_last (physically ``[Cout][R][S][Cin]``) plus a bf16 shadow.

Every op here launches our kernels; there is no eager-PyTorch fallback on a CUDA tensor.  The same
math in plain PyTorch lives in ``ops/reference.py`` and is what the tests compare against (and what
CPU tensors use, so the model classes stay importable/runnable for plumbing tests without a GPU).

Reference call sites: every layer of torchvision ``ResNet(Bottleneck,[3,4,6,3])`` used by
``R/03 模型并行/03_model_parallel.ipynb#cell4-5,12`` and the loss/optimizer calls of
``R/02 DDP基本概念/ddp_gpus.py:37-42``.
"""
from __future__ import annotations

import ctypes
import functools
import os
import weakref
from typing import Optional, Tuple

import torch

from .. import _lib
from . import reference as ref

CL = torch.channels_last


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _on_device(fn):
    """Run ``fn`` with its first argument's GPU as the current device.  Kernels launch on the CURRENT device and
    stream: a module living on cuda:1 that is called while cuda:0 is current (a model-parallel forward) would
    otherwise run on GPU 0's SMs and reach its own tensors over NVLink - or fault where peer access is off."""
    @functools.wraps(fn)
    def wrapper(x, *args, **kwargs):
        if isinstance(x, torch.Tensor) and x.is_cuda and x.device.index != torch.cuda.current_device():
            with torch.cuda.device(x.device):
                return fn(x, *args, **kwargs)
        return fn(x, *args, **kwargs)
    return wrapper


# ------------------------------------------------------------------------------------------------
# gradient sink: lets a data-parallel wrapper receive parameter gradients where it wants them
# ------------------------------------------------------------------------------------------------
# A sink has two methods:
#   targets(params) -> list of fp32 tensors (same element order as each parameter) that the kernels may
#                      ACCUMULATE into directly, or None to decline (then autograd's AccumulateGrad runs as usual);
#   done(params)    -> called once the kernels writing those targets have been issued on the current stream.
# ``DistributedDataParallel`` installs itself: weight gradients are then reduced (red.global.add) straight into
# the gradient-bucket views, no temporary and no per-parameter add kernel.
_GRAD_SINK = None      # weak reference: the sink's lifetime stays with whoever built it


def set_grad_sink(sink) -> None:
    global _GRAD_SINK
    _GRAD_SINK = weakref.ref(sink) if sink is not None else None


def grad_sink():
    return _GRAD_SINK() if _GRAD_SINK is not None else None


def _empty_act(n: int, c: int, h: int, w: int, device, dtype=torch.bfloat16) -> torch.Tensor:
    return torch.empty((n, c, h, w), device=device, dtype=dtype, memory_format=CL)


def is_nhwc_bf16(x: torch.Tensor) -> bool:
    return x.dtype == torch.bfloat16 and x.dim() == 4 and x.is_contiguous(memory_format=CL)


@_on_device
def as_nhwc_bf16(x: torch.Tensor) -> torch.Tensor:
    """Bring any 4-D activation to bf16 channels_last (no copy if already there)."""
    if is_nhwc_bf16(x):
        return x
    if x.is_cuda and x.dtype == torch.float32 and x.is_contiguous():
        n, c, h, w = x.shape
        out = _empty_act(n, c, h, w, x.device)
        _lib.call("ptd_nchw_f32_to_nhwc_bf16", x, out, n, c, h * w, _stream())
        return out
    return x.to(dtype=torch.bfloat16).contiguous(memory_format=CL)


def out_hw(h: int, w: int, r: int, s: int, stride: int, pad: int) -> Tuple[int, int]:
    return (h + 2 * pad - r) // stride + 1, (w + 2 * pad - s) // stride + 1


# ------------------------------------------------------------------------------------------------
# raw kernel wrappers (no autograd)
# ------------------------------------------------------------------------------------------------
def conv_fprop_raw(x: torch.Tensor, w_bf16: torch.Tensor, stride: int, pad: int, want_stats: bool):
    n, cin, h, w = x.shape
    cout, _, r, s = w_bf16.shape
    oh, ow = out_hw(h, w, r, s, stride, pad)
    y = _empty_act(n, cout, oh, ow, x.device)
    stats = torch.zeros((2, cout), device=x.device, dtype=torch.float32) if want_stats else None
    _lib.call("ptd_conv_fprop", x, w_bf16, y, stats[0] if want_stats else None, stats[1] if want_stats else None,
              n, h, w, cin, cout, r, s, stride, pad, _stream())
    return y, stats


def conv_dgrad_raw(dy: torch.Tensor, w_bf16: torch.Tensor, x_shape, stride: int, pad: int,
                   addend: Optional[torch.Tensor] = None) -> torch.Tensor:
    n, cin, h, w = x_shape
    cout, _, r, s = w_bf16.shape
    dx = _empty_act(n, cin, h, w, dy.device)
    _lib.call("ptd_conv_dgrad", dy, w_bf16, dx, addend, n, h, w, cin, cout, r, s, stride, pad, _stream())
    return dx


def conv_wgrad_raw(x: torch.Tensor, dy: torch.Tensor, w_shape, stride: int, pad: int,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    n, cin, h, w = x.shape
    cout, _, r, s = w_shape
    if out is None:
        out = torch.empty(w_shape, device=x.device, dtype=torch.float32, memory_format=CL).zero_()
    _lib.call("ptd_conv_wgrad", x, dy, out, n, h, w, cin, cout, r, s, stride, pad, _stream())
    return out


def bn_apply_raw(y, stats, gamma, beta, running_mean, running_var, eps, momentum, relu, residual=None,
                 want_mask: bool = False):
    """Returns ``(z, saved)``, or ``(z, saved, mask)`` with ``want_mask``: ``mask`` is the packed ReLU bit tensor
    (uint8 ``[M, C/8]``) that :func:`bn_backward_raw` reads instead of ``z`` (``None`` when ``relu`` is off)."""
    n, c, h, w = y.shape
    z = torch.empty_like(y, memory_format=CL)
    saved = torch.empty((2, c), device=y.device, dtype=torch.float32)
    mask = torch.empty((n * h * w, c // 8), device=y.device, dtype=torch.uint8) if (want_mask and relu) else None
    _lib.call("ptd_bn_apply_mask", y, residual, z, mask, stats[0], stats[1], gamma, beta, running_mean, running_var,
              saved[0], saved[1], n * h * w, c, float(eps), float(momentum), int(relu), _stream())
    return (z, saved, mask) if want_mask else (z, saved)


def bn_stats_raw(y: torch.Tensor) -> torch.Tensor:
    n, c, h, w = y.shape
    stats = torch.zeros((2, c), device=y.device, dtype=torch.float32)
    _lib.call("ptd_bn_stats", y, stats[0], stats[1], n * h * w, c, _stream())
    return stats


def bn_backward_raw(dz, z, y, saved, gamma, relu: bool, want_gout: bool, mask: Optional[torch.Tensor] = None):
    """``mask`` (from ``bn_apply_raw(..., want_mask=True)``) replaces ``z`` as the source of the ReLU mask."""
    n, c, h, w = y.shape
    m = n * h * w
    sums = torch.zeros((2, c), device=y.device, dtype=torch.float32)   # [0] = dbeta, [1] = dgamma
    mode, zm = (2, mask) if (relu and mask is not None) else (int(relu), z)
    _lib.call("ptd_bn_bwd_reduce", dz, zm, y, saved[0], saved[1], sums[0], sums[1], m, c, mode, _stream())
    dy = torch.empty_like(y, memory_format=CL)
    gout = torch.empty_like(y, memory_format=CL) if want_gout else None
    _lib.call("ptd_bn_bwd_apply", dz, zm, y, saved[0], saved[1], gamma, sums[0], sums[1], dy, gout, m, c, mode,
              _stream())
    return dy, gout, sums[1], sums[0]


# ------------------------------------------------------------------------------------------------
# conv + batch-norm (+ residual) (+ ReLU), training mode - the workhorse
# ------------------------------------------------------------------------------------------------
class _ConvBnAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, w_bf16, gamma, beta, running_mean, running_var, residual, stride, pad, eps,
                momentum, relu, stem):
        x = x if stem else as_nhwc_bf16(x)
        if stem:
            y, stats = stem_fprop_raw(x, w_bf16)
        else:
            y, stats = conv_fprop_raw(x, w_bf16, stride, pad, True)
        z, saved, mask = bn_apply_raw(y, stats, gamma, beta, running_mean, running_var, eps, momentum, relu, residual,
                                      want_mask=True)
        ctx.save_for_backward(x, w_bf16, y, mask, saved, gamma)     # the bit mask stands in for z
        ctx.cfg = (stride, pad, relu, stem, residual is not None, tuple(weight.shape))
        return z

    @staticmethod
    def backward(ctx, dz):
        x, w_bf16, y, mask, saved, gamma = ctx.saved_tensors
        stride, pad, relu, stem, has_res, w_shape = ctx.cfg
        dz = as_nhwc_bf16(dz)
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        dy, gout, dgamma, dbeta = bn_backward_raw(dz, None, y, saved, gamma, relu,
                                                  has_res and ctx.needs_input_grad[7], mask=mask)
        dx = dw = None
        if stem:
            if need_w:
                dw = stem_wgrad_raw(x, dy, w_shape)
        else:
            if need_x:
                dx = conv_dgrad_raw(dy, w_bf16, x.shape, stride, pad)
            if need_w:
                dw = conv_wgrad_raw(x, dy, w_shape, stride, pad)
        return dx, dw, None, dgamma, dbeta, None, None, gout, None, None, None, None, None, None


@_on_device
def conv_bn_act(x, weight, w_bf16, gamma, beta, running_mean, running_var, residual=None, stride=1, pad=0,
                eps=1e-5, momentum=0.1, relu=True, stem=False):
    if not x.is_cuda:
        return ref.conv_bn_act(x, weight, gamma, beta, running_mean, running_var, residual, stride, pad, eps,
                               momentum, relu)
    return _ConvBnAct.apply(x, weight, w_bf16, gamma, beta, running_mean, running_var, residual, stride, pad, eps,
                            momentum, relu, stem)


# ---- conv only / bn only (standalone modules, e.g. inside a user-built nn.Sequential) -------------
class _Conv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, w_bf16, stride, pad):
        x = as_nhwc_bf16(x)
        y, _ = conv_fprop_raw(x, w_bf16, stride, pad, False)
        ctx.save_for_backward(x, w_bf16)
        ctx.cfg = (stride, pad, tuple(weight.shape))
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w_bf16 = ctx.saved_tensors
        stride, pad, w_shape = ctx.cfg
        dy = as_nhwc_bf16(dy)
        dx = conv_dgrad_raw(dy, w_bf16, x.shape, stride, pad) if ctx.needs_input_grad[0] else None
        dw = conv_wgrad_raw(x, dy, w_shape, stride, pad) if ctx.needs_input_grad[1] else None
        return dx, dw, None, None, None


@_on_device
def conv2d(x, weight, w_bf16, stride=1, pad=0):
    if not x.is_cuda:
        return ref.conv2d(x, weight, stride, pad)
    return _Conv.apply(x, weight, w_bf16, stride, pad)


class _BnAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, gamma, beta, running_mean, running_var, residual, eps, momentum, relu):
        y = as_nhwc_bf16(y)
        stats = bn_stats_raw(y)
        z, saved, mask = bn_apply_raw(y, stats, gamma, beta, running_mean, running_var, eps, momentum, relu, residual,
                                      want_mask=True)
        ctx.save_for_backward(y, mask, saved, gamma)
        ctx.cfg = (relu, residual is not None)
        return z

    @staticmethod
    def backward(ctx, dz):
        y, mask, saved, gamma = ctx.saved_tensors
        relu, has_res = ctx.cfg
        dz = as_nhwc_bf16(dz)
        dy, gout, dgamma, dbeta = bn_backward_raw(dz, None, y, saved, gamma, relu,
                                                  has_res and ctx.needs_input_grad[5], mask=mask)
        return dy, dgamma, dbeta, None, None, gout, None, None, None


@_on_device
def batch_norm_act(y, gamma, beta, running_mean, running_var, residual=None, eps=1e-5, momentum=0.1, relu=False):
    if not y.is_cuda:
        return ref.batch_norm_act(y, gamma, beta, running_mean, running_var, residual, eps, momentum, relu)
    return _BnAct.apply(y, gamma, beta, running_mean, running_var, residual, eps, momentum, relu)


@_on_device
def affine_act(y, scale, shift, residual=None, relu=False):
    """Inference-mode BN (per-channel affine) + residual + ReLU; no autograd (eval only)."""
    if not y.is_cuda:
        return ref.affine_act(y, scale, shift, residual, relu)
    y = as_nhwc_bf16(y)
    n, c, h, w = y.shape
    z = torch.empty_like(y, memory_format=CL)
    _lib.call("ptd_affine_act", y, residual, z, scale, shift, n * h * w, c, int(relu), _stream())
    return z


# ------------------------------------------------------------------------------------------------
# stem helpers
# ------------------------------------------------------------------------------------------------
@_on_device
def to_nhwc4(x: torch.Tensor) -> torch.Tensor:
    """fp32 NCHW image batch [N,3,H,W] -> bf16 NHWC with a zero 4th channel (stem input, any H x W)."""
    n, c, h, w = x.shape
    assert c == 3, "stem expects 3 input channels"
    x = x.contiguous().float()
    out = _empty_act(n, 4, h, w, x.device)
    _lib.call("ptd_nchw3_f32_to_nhwc4_bf16", x, out, n, h, w, _stream())
    return out


def stem_s2d_ok(h: int, w: int) -> bool:
    """The TMA stem works on whole 8 x 16 patches of output pixels."""
    return h % 16 == 0 and w % 32 == 0 and os.environ.get("PTDB200_STEM_S2D", "1") != "0"


@_on_device
def to_s2d(x: torch.Tensor) -> torch.Tensor:
    """fp32 NCHW image batch [N,3,H,W] -> bf16 space-to-depth stem input, logical shape [N,16,H/2+3,W/2+3]
    (channels_last): pixel (hs, ws) = the 2x2 block at (2hs-3, 2ws-3) of the zero-padded image, 12 of 16 channels
    used.  On it the 7x7/2 stem is a 4x4/1 convolution whose im2col rows are contiguous (``csrc/igemm.cu``)."""
    n, c, h, w = x.shape
    assert c == 3 and stem_s2d_ok(h, w)
    x = x.contiguous().float()
    out = _empty_act(n, 16, h // 2 + 3, w // 2 + 3, x.device)
    _lib.call("ptd_nchw3_f32_to_s2d_bf16", x, out, n, h, w, _stream())
    return out


def stem_input(x: torch.Tensor) -> torch.Tensor:
    """The stem's input format for this image size: space-to-depth when the size allows, NHWC4 otherwise."""
    return to_s2d(x) if stem_s2d_ok(x.shape[2], x.shape[3]) else to_nhwc4(x)


@_on_device
def stem_pack(weight: torch.Tensor) -> torch.Tensor:
    """bf16 [2][64][256]: [0] = [64][8][8][4] for the NHWC4 stem, [1] = [64][4][4][16] for the space-to-depth stem."""
    wp = torch.empty((2, 64, 256), device=weight.device, dtype=torch.bfloat16)
    sk, sc, sr, ss = weight.stride()
    _lib.call("ptd_stem_pack_weight", weight, wp[0], sk, sc, sr, ss, _stream())
    _lib.call("ptd_stem_pack_weight_s2d", weight, wp[1], sk, sc, sr, ss, _stream())
    return wp


@_on_device
def stem_fprop_raw(xin: torch.Tensor, wp: torch.Tensor):
    """``xin`` from :func:`stem_input` (16 channels = space-to-depth, 4 = NHWC4), ``wp`` from :func:`stem_pack`."""
    n, c, h, w = xin.shape
    stats = torch.zeros((2, 64), device=xin.device, dtype=torch.float32)
    if c == 16:
        oh, ow = h - 3, w - 3
        y = _empty_act(n, 64, oh, ow, xin.device)
        _lib.call("ptd_stem_s2d_fprop", xin, wp[1], y, stats[0], stats[1], n, 2 * oh, 2 * ow, _stream())
    else:
        oh, ow = out_hw(h, w, 7, 7, 2, 3)
        y = _empty_act(n, 64, oh, ow, xin.device)
        _lib.call("ptd_stem_fprop", xin, wp[0], y, stats[0], stats[1], n, h, w, _stream())
    return y, stats


def stem_wgrad_raw(xin: torch.Tensor, dy: torch.Tensor, w_shape) -> torch.Tensor:
    n, c, h, w = xin.shape
    dwp = torch.zeros((64, 256), device=xin.device, dtype=torch.float32)
    dw = torch.empty(w_shape, device=xin.device, dtype=torch.float32, memory_format=CL)
    sk, sc, sr, ss = dw.stride()
    if c == 16:
        _lib.call("ptd_stem_s2d_wgrad", xin, dy, dwp, n, 2 * (h - 3), 2 * (w - 3), _stream())
        _lib.call("ptd_stem_unpack_wgrad_s2d", dwp, dw, sk, sc, sr, ss, 0, _stream())
    else:
        _lib.call("ptd_stem_wgrad", xin, dy, dwp, n, h, w, _stream())
        _lib.call("ptd_stem_unpack_wgrad", dwp, dw, sk, sc, sr, ss, 0, _stream())
    return dw


class _StemBnReluPool(torch.autograd.Function):
    """stem conv (7x7/2) -> BN (batch stats) -> ReLU -> maxpool 3x3/2 with the BN/ReLU/pool part in ONE kernel
    forward (``ptd_bn_relu_maxpool_fwd``) and two backward (``ptd_bn_relu_maxpool_bwd``): the 112x112 post-BN
    activation and its gradient are never materialised.  ``xin`` comes from :func:`stem_input`."""

    @staticmethod
    def forward(ctx, xin, weight, wp, gamma, beta, running_mean, running_var, eps, momentum):
        y, stats = stem_fprop_raw(xin, wp)
        n, c, h, w = y.shape
        p = _empty_act(n, c, h // 2, w // 2, y.device)
        idx = torch.empty((n, h // 2, w // 2, c), device=y.device, dtype=torch.uint8)
        saved = torch.empty((2, c), device=y.device, dtype=torch.float32)
        _lib.call("ptd_bn_relu_maxpool_fwd", y, p, idx, stats[0], stats[1], gamma, beta, running_mean, running_var,
                  saved[0], saved[1], n, h, w, c, float(eps), float(momentum), _stream())
        ctx.save_for_backward(xin, y, idx, saved, gamma)
        ctx.w_shape = tuple(weight.shape)
        return p

    @staticmethod
    def backward(ctx, dp):
        xin, y, idx, saved, gamma = ctx.saved_tensors
        dp = as_nhwc_bf16(dp)
        n, c, h, w = y.shape
        sums = torch.zeros((2, c), device=y.device, dtype=torch.float32)        # [dbeta | dgamma]
        dy = torch.empty_like(y, memory_format=CL)
        _lib.call("ptd_bn_relu_maxpool_bwd", dp, idx, y, saved[0], saved[1], gamma, sums, dy, n, h, w, c, _stream())
        dw = stem_wgrad_raw(xin, dy, ctx.w_shape) if ctx.needs_input_grad[1] else None
        return None, dw, None, sums[1], sums[0], None, None, None, None


@_on_device
def stem_bn_relu_pool(xin, weight, wp, gamma, beta, running_mean, running_var, eps=1e-5, momentum=0.1):
    """Training-mode stem + BN + ReLU + maxpool on a CUDA stem input whose conv output has even height and width."""
    return _StemBnReluPool.apply(xin, weight, wp, gamma, beta, running_mean, running_var, eps, momentum)


# ------------------------------------------------------------------------------------------------
# pooling
# ------------------------------------------------------------------------------------------------
class _MaxPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = as_nhwc_bf16(x)
        n, c, h, w = x.shape
        oh, ow = out_hw(h, w, 3, 3, 2, 1)
        y = _empty_act(n, c, oh, ow, x.device)
        idx = _empty_act(n, c, oh, ow, x.device, dtype=torch.uint8)
        _lib.call("ptd_maxpool3x3s2_fwd", x, y, idx, n, h, w, c, _stream())
        ctx.save_for_backward(idx)
        ctx.shape = (n, c, h, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        n, c, h, w = ctx.shape
        dy = as_nhwc_bf16(dy)
        dx = _empty_act(n, c, h, w, dy.device)
        _lib.call("ptd_maxpool3x3s2_bwd", dy, idx, dx, n, h, w, c, _stream())
        return dx


@_on_device
def max_pool_3x3_s2(x):
    if not x.is_cuda:
        return ref.max_pool_3x3_s2(x)
    return _MaxPool.apply(x)


class _AvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = as_nhwc_bf16(x)
        n, c, h, w = x.shape
        y = torch.empty((n, c), device=x.device, dtype=torch.bfloat16)
        _lib.call("ptd_avgpool_fwd", x, y, n, h * w, c, _stream())
        ctx.shape = (n, c, h, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        n, c, h, w = ctx.shape
        dy = dy.to(torch.bfloat16).contiguous()
        dx = _empty_act(n, c, h, w, dy.device)
        _lib.call("ptd_avgpool_bwd", dy, dx, n, h * w, c, _stream())
        return dx


@_on_device
def global_avg_pool(x):
    """[N,C,H,W] -> [N,C] (bf16), i.e. AdaptiveAvgPool2d(1) + flatten."""
    if not x.is_cuda:
        return ref.global_avg_pool(x)
    return _AvgPool.apply(x)


# ------------------------------------------------------------------------------------------------
# classifier, loss
# ------------------------------------------------------------------------------------------------
class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, weight, w_bf16, bias, out_box=None):
        a = a.to(torch.bfloat16).contiguous()
        m, k = a.shape
        n = weight.shape[0]
        # out_box: a one-element list holding a preallocated fp32 [m, n] row block (e.g. the rows of a micro-batch inside
        # the pipeline's output tensor) that the kernel writes in place of a fresh tensor.  Boxed so that autograd does
        # not see it as an input of this node.
        if out_box is not None:
            out = out_box[0]
            assert out.shape == (m, n) and out.dtype == torch.float32 and out.is_contiguous() and out.device == a.device
        else:
            out = torch.empty((m, n), device=a.device, dtype=torch.float32)
        _lib.call("ptd_linear_fwd", a, w_bf16, bias, out, m, n, k, _stream())
        ctx.save_for_backward(a, w_bf16)
        ctx.has_bias = bias is not None
        return out

    @staticmethod
    def backward(ctx, dout):
        a, w_bf16 = ctx.saved_tensors
        m, k = a.shape
        n = w_bf16.shape[0]
        d16 = dout.to(torch.bfloat16).contiguous()
        da = dw = db = None
        if ctx.needs_input_grad[0]:
            da = torch.empty((m, k), device=a.device, dtype=torch.bfloat16)
            _lib.call("ptd_linear_dgrad", d16, w_bf16, da, m, n, k, _stream())
        if ctx.needs_input_grad[1]:
            dw = torch.zeros((n, k), device=a.device, dtype=torch.float32)
            _lib.call("ptd_linear_wgrad", a, d16, dw, m, n, k, _stream())
        if ctx.has_bias and ctx.needs_input_grad[3]:
            db = dout.float().sum(0)
        return da, dw, None, db, None


@_on_device
def linear(a, weight, w_bf16, bias, out=None):
    """``out``: optional preallocated fp32 ``[rows, out_features]`` block the result is written into (CUDA only)."""
    if not a.is_cuda:
        res = ref.linear(a, weight, bias)
        if out is not None:
            out.copy_(res.detach())
        return res
    return _Linear.apply(a, weight, w_bf16, bias, None if out is None else [out])


class _GatherRows(torch.autograd.Function):
    """The pieces ARE consecutive row blocks of ``buf`` already (each was written there by its producer): returns
    ``buf`` as one tensor without a concatenation pass; the backward hands every piece its rows of the gradient."""

    @staticmethod
    def forward(ctx, buf_box, *pieces):
        buf = buf_box[0]
        ctx.sizes = [p.size(0) for p in pieces]
        row = 0
        for p in pieces:
            if p.data_ptr() != buf[row:row + p.size(0)].data_ptr():      # a producer that could not write in place
                buf[row:row + p.size(0)].copy_(p)
            row += p.size(0)
        return buf.view_as(buf)

    @staticmethod
    def backward(ctx, g):
        return (None,) + tuple(g.split(ctx.sizes, dim=0))


def gather_rows(buf: torch.Tensor, pieces):
    """``torch.cat(pieces)`` for pieces that already live in consecutive row blocks of ``buf``."""
    return _GatherRows.apply([buf], *pieces)


class _SoftmaxCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        logits = logits.float().contiguous()
        rows, c = logits.shape
        loss = torch.zeros((), device=logits.device, dtype=torch.float32)
        dlogits = torch.empty_like(logits)
        soft = target.is_floating_point()
        tgt = target.float().contiguous() if soft else target.long().contiguous()
        _lib.call("ptd_softmax_ce", logits, tgt if soft else None, None if soft else tgt, loss, dlogits, 0, rows, c,
                  c, 1.0 / rows, _stream())
        ctx.save_for_backward(dlogits)
        return loss

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        return dlogits * g, None


@_on_device
def cross_entropy(logits, target):
    """Mean softmax cross-entropy; ``target`` is class indices [B] or probabilities [B, C]."""
    if not logits.is_cuda:
        return ref.cross_entropy(logits, target)
    return _SoftmaxCE.apply(logits, target)


# ------------------------------------------------------------------------------------------------
# tutorial-sized dense layers (fp32, CUDA cores)
# ------------------------------------------------------------------------------------------------
class _SmallLinear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, relu):
        x = x.float().contiguous()
        bsz, k = x.shape
        n = w.shape[0]
        y = torch.empty((bsz, n), device=x.device, dtype=torch.float32)
        _lib.call("ptd_small_linear_fwd", x, w.contiguous(), b, y, bsz, k, n, int(relu), _stream())
        ctx.save_for_backward(x, w, y)
        ctx.relu = relu
        ctx.has_bias = b is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        bsz, k = x.shape
        n = w.shape[0]
        dy = dy.float().contiguous()
        if ctx.relu:
            dy = dy * (y > 0)
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        dw = torch.zeros_like(w, memory_format=torch.contiguous_format) if ctx.needs_input_grad[1] else None
        db = torch.zeros(n, device=x.device, dtype=torch.float32) if (ctx.has_bias and dw is not None) else None
        _lib.call("ptd_small_linear_bwd", x, w.contiguous(), dy, dx, dw, db, bsz, k, n, _stream())
        return dx, dw, db, None


@_on_device
def small_linear(x, w, b, relu=False):
    if not x.is_cuda:
        return ref.small_linear(x, w, b, relu)
    return _SmallLinear.apply(x, w, b, relu)


# ------------------------------------------------------------------------------------------------
# a whole residual bottleneck as ONE autograd node
# ------------------------------------------------------------------------------------------------
def _ptr4(ts):
    arr = (ctypes.c_void_p * 4)()
    for i, t in enumerate(ts):
        arr[i] = t.data_ptr() if t is not None else None
    return arr


class _Bottleneck(torch.autograd.Function):
    """conv1x1-BN-ReLU, conv3x3-BN-ReLU, conv1x1-BN (+identity or +downsample conv-BN) -ReLU as ONE autograd node
    whose forward and backward are ONE native call each (``csrc/block.cu``).

    One node instead of four keeps the Python/autograd overhead per block small and lets the backward fuse what
    autograd would do with separate kernels: the gradient of the identity branch (the masked incoming gradient,
    emitted by the BN backward of the last unit) is added inside the epilogue of the first convolution's
    data-gradient kernel, the shortcut convolution's data gradient accumulates in place, and every buffer of the
    block comes from two allocations per direction.

    ``tensors`` = for each unit (conv1, conv2, conv3[, downsample]): weight, w_bf16, gamma, beta, running_mean,
    running_var.  ``cfg`` = (stride, eps, momentum, has_downsample).
    """

    @staticmethod
    def forward(ctx, x, cfg, *tensors):
        stride, eps, momentum, has_ds = cfg
        x = as_nhwc_bf16(x)
        nu = 4 if has_ds else 3
        units = [tensors[i * 6:(i + 1) * 6] for i in range(nu)]
        n, inplanes, h, w = x.shape
        planes = units[0][0].shape[0]
        oh, ow = out_hw(h, w, 3, 3, stride, 1)
        shapes = [(n, planes, h, w), (n, planes, oh, ow), (n, planes * 4, oh, ow), (n, planes * 4, oh, ow)][:nu]
        couts = [s_[1] for s_ in shapes]
        numels = [s_[0] * s_[1] * s_[2] * s_[3] for s_ in shapes]
        # y tensors (kept for the backward) and z tensors in one allocation each
        ybuf = torch.empty(sum(numels), device=x.device, dtype=torch.bfloat16)
        zbuf = torch.empty(sum(numels), device=x.device, dtype=torch.bfloat16)
        ys, zs, o = [], [], 0
        for shp, nel in zip(shapes, numels):
            nn_, c_, h_, w_ = shp
            strides = (h_ * w_ * c_, 1, w_ * c_, c_)
            ys.append(torch.as_strided(ybuf, shp, strides, o))
            zs.append(torch.as_strided(zbuf, shp, strides, o))
            o += nel
        fbuf = torch.empty(4 * sum(couts), device=x.device, dtype=torch.float32)   # [stats | saved]
        stats, saved = fbuf[:2 * sum(couts)], fbuf[2 * sum(couts):]
        d = _lib.BottleneckFwd()
        d.N, d.H, d.W, d.inplanes, d.planes, d.stride, d.has_ds = n, h, w, inplanes, planes, stride, int(has_ds)
        d.eps, d.momentum = float(eps), float(momentum)
        d.x = x.data_ptr()
        d.w = _ptr4([u[1] for u in units])
        d.gamma = _ptr4([u[2] for u in units])
        d.beta = _ptr4([u[3] for u in units])
        d.running_mean = _ptr4([u[4] for u in units])
        d.running_var = _ptr4([u[5] for u in units])
        d.y, d.z = _ptr4(ys), _ptr4(zs)
        d.stats, d.saved = stats.data_ptr(), saved.data_ptr()
        # ReLU bit masks of the three main-path units (1 bit per element): what the BN backward reads instead of z
        mbuf = torch.empty(sum(numels[:3]) // 8, device=x.device, dtype=torch.uint8)
        ma, o = (ctypes.c_void_p * 4)(), 0
        for i in range(3):
            ma[i] = mbuf.data_ptr() + o
            o += numels[i] // 8
        d.mask = ma
        _lib.call("ptd_bottleneck_fwd", d, _stream())
        ctx.cfg = cfg
        ctx.geom = (n, h, w, inplanes, planes, shapes)
        ctx.w_shapes = [tuple(u[0].shape) for u in units]
        # the parameter objects themselves (weight, gamma, beta per unit): what a gradient sink is asked about
        ctx.params = [t for u in units for t in (u[0], u[2], u[3])]
        ctx.save_for_backward(x, ybuf, zbuf, saved, mbuf, *[u[1] for u in units], *[u[2] for u in units])
        return zs[2]

    @staticmethod
    def backward(ctx, dz):
        stride, eps, momentum, has_ds = ctx.cfg
        nu = 4 if has_ds else 3
        n, h, w, inplanes, planes, shapes = ctx.geom
        sv_t = ctx.saved_tensors
        x, ybuf, zbuf, saved, mbuf = sv_t[:5]
        wbs = sv_t[5:5 + nu]
        gammas = sv_t[5 + nu:5 + 2 * nu]
        w_shapes = ctx.w_shapes
        dz = as_nhwc_bf16(dz)
        dev = dz.device
        couts = [s_[1] for s_ in shapes]
        numels = [s_[0] * s_[1] * s_[2] * s_[3] for s_ in shapes]
        esz = 2
        ys_ptr, zs_ptr, o = [], [], 0
        for nel in numels:
            ys_ptr.append(ybuf.data_ptr() + o * esz)
            zs_ptr.append(zbuf.data_ptr() + o * esz)
            o += nel
        # scratch: dy per unit, g_id (block-output sized), dz2 (z[1] sized), dz1 (z[0] sized); output dx
        x_numel = x.numel()
        sizes = numels + [numels[2], numels[1], numels[0], x_numel]
        abuf = torch.empty(sum(sizes), device=dev, dtype=torch.bfloat16)
        offs, o = [], 0
        for sz in sizes:
            offs.append(o)
            o += sz
        base = abuf.data_ptr()
        w_numels = [s_[0] * s_[1] * s_[2] * s_[3] for s_ in w_shapes]
        # A gradient sink (DDP) may take every parameter gradient of the block where it lives: the weight-gradient
        # kernels then accumulate into its tensors and dbeta / dgamma are added there by the same native call.
        sink = grad_sink()
        direct = sink.targets(ctx.params) if sink is not None else None
        if direct is not None:
            gbuf = torch.empty(2 * sum(couts), device=dev, dtype=torch.float32)
            sums, dw_all = gbuf, None
            dws = [direct[3 * i] for i in range(nu)]
        else:
            gbuf = torch.empty(2 * sum(couts) + sum(w_numels), device=dev, dtype=torch.float32)   # [sums | dw_all]
            sums = gbuf[:2 * sum(couts)]
            dw_all = gbuf[2 * sum(couts):]
            dws, o = [], 0
            for shp, nel in zip(w_shapes, w_numels):
                co, ci, r, s_ = shp
                # as_strided offsets are absolute in the storage: dw_all starts after the sums
                dws.append(torch.as_strided(gbuf, shp, (ci * r * s_, 1, s_ * ci, ci), 2 * sum(couts) + o))   # channels_last view
                o += nel
        d = _lib.BottleneckBwd()
        d.N, d.H, d.W, d.inplanes, d.planes, d.stride, d.has_ds = n, h, w, inplanes, planes, stride, int(has_ds)
        d.x, d.dz = x.data_ptr(), dz.data_ptr()
        d.w, d.gamma = _ptr4(wbs), _ptr4(gammas)
        ya, za = (ctypes.c_void_p * 4)(), (ctypes.c_void_p * 4)()
        for i in range(nu):
            ya[i], za[i] = ys_ptr[i], zs_ptr[i]
        d.y, d.z = ya, za
        ma, o = (ctypes.c_void_p * 4)(), 0
        for i in range(3):
            ma[i] = mbuf.data_ptr() + o
            o += numels[i] // 8
        d.mask = ma
        d.saved, d.sums = saved.data_ptr(), sums.data_ptr()
        if direct is not None:
            d.dw_all, d.dw_all_bytes = None, 0
            gd = (ctypes.c_void_p * 8)()
            for i in range(nu):
                gd[2 * i] = direct[3 * i + 2].data_ptr()        # dbeta -> bias.grad
                gd[2 * i + 1] = direct[3 * i + 1].data_ptr()    # dgamma -> weight.grad of the BatchNorm
            d.gdst = gd
        else:
            d.dw_all, d.dw_all_bytes = dw_all.data_ptr(), dw_all.numel() * 4
        d.dw = _ptr4(dws)
        dya = (ctypes.c_void_p * 4)()
        for i in range(nu):
            dya[i] = base + offs[i] * esz
        d.dy = dya
        d.g_id = base + offs[nu] * esz
        d.dz2 = base + offs[nu + 1] * esz
        d.dz1 = base + offs[nu + 2] * esz
        d.dx = base + offs[nu + 3] * esz
        _lib.call("ptd_bottleneck_bwd", d, _stream())
        xs = x.shape
        dx = torch.as_strided(abuf, xs, (xs[2] * xs[3] * xs[1], 1, xs[3] * xs[1], xs[1]), offs[nu + 3])
        if direct is not None:
            sink.done(ctx.params)
            return (dx, None) + (None,) * (6 * nu)
        so, grads = 0, [dx, None]
        for i in range(nu):
            c = couts[i]
            grads += [dws[i], None, sums[so + c:so + 2 * c], sums[so:so + c], None, None]
            so += 2 * c
        return tuple(grads)


@_on_device
def bottleneck(x, units, stride: int, eps: float, momentum: float):
    """``units``: list of (weight, w_bf16, gamma, beta, running_mean, running_var) for conv1, conv2, conv3 and
    optionally the downsample branch.  CUDA, training mode only."""
    flat = [t for u in units for t in u]
    return _Bottleneck.apply(x, (stride, eps, momentum, len(units) == 4), *flat)
