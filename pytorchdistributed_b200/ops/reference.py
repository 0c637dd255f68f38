"""Plain-PyTorch statement of every op in ``functional.py``.

Two uses: (1) the oracle the GPU numerics tests compare our kernels against (run in fp32), and
(2) the path CPU tensors take, so models, wrappers and scripts can be exercised on a box without
a GPU (gloo plumbing tests).  Nothing here is on a CUDA hot path.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def conv2d(x, weight, stride=1, pad=0):
    return F.conv2d(x, weight.to(x.dtype), None, stride, pad)


def batch_norm_act(y, gamma, beta, running_mean, running_var, residual=None, eps=1e-5, momentum=0.1, relu=False):
    z = F.batch_norm(y.float(), running_mean, running_var, gamma, beta, True, momentum, eps).to(y.dtype)
    if residual is not None:
        z = z + residual
    return F.relu(z) if relu else z


def conv_bn_act(x, weight, gamma, beta, running_mean, running_var, residual=None, stride=1, pad=0, eps=1e-5,
                momentum=0.1, relu=True):
    y = conv2d(x, weight, stride, pad)
    return batch_norm_act(y, gamma, beta, running_mean, running_var, residual, eps, momentum, relu)


def affine_act(y, scale, shift, residual=None, relu=False):
    z = (y.float() * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)).to(y.dtype)
    if residual is not None:
        z = z + residual
    return F.relu(z) if relu else z


def max_pool_3x3_s2(x):
    return F.max_pool2d(x, 3, 2, 1)


def global_avg_pool(x):
    return x.float().mean((2, 3)).to(x.dtype)


def linear(a, weight, bias):
    return F.linear(a.float(), weight.float(), bias)


def cross_entropy(logits, target):
    return F.cross_entropy(logits.float(), target)


def small_linear(x, w, b, relu=False):
    y = F.linear(x, w, b)
    return F.relu(y) if relu else y
