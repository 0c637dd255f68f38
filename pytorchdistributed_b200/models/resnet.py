"""ResNet (Bottleneck) on the sm_100a kernels, ``state_dict``-compatible with torchvision.

What the reference builds: ``ResNet(Bottleneck, [3, 4, 6, 3], num_classes=1000)``
(``R/03 模型并行/03_model_parallel.ipynb#cell4``, 25,557,032 parameters) with the stride on the 3x3
convolution (torchvision "v1.5").  Attribute names (``conv1, bn1, relu, maxpool, layer1..4,
avgpool, fc``) match so the reference's way of carving the model into ``nn.Sequential`` stages
(``#cell5``) carries over; ``stage_modules()`` gives the same two stages with the fused kernels.

Execution differs from torchvision's module-by-module walk: each conv+BN(+add)(+ReLU) group is
one fused autograd node (conv epilogue produces the BN batch statistics; BN apply, residual add
and ReLU are one pass), activations stay NHWC bf16 end to end.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import os

import torch
from torch import nn

from ..ops import functional as F_
from .layers import BatchNorm2d, Conv2d, GlobalAvgPool, Linear, MaxPool2d, conv_bn_act


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None) -> None:
        super().__init__()
        width = planes
        self.conv1 = Conv2d(inplanes, width, 1)
        self.bn1 = BatchNorm2d(width)
        self.conv2 = Conv2d(width, width, 3, stride, 1)
        self.bn2 = BatchNorm2d(width)
        self.conv3 = Conv2d(width, planes * self.expansion, 1)
        self.bn3 = BatchNorm2d(planes * self.expansion)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def _units(self):
        pairs = [(self.conv1, self.bn1), (self.conv2, self.bn2), (self.conv3, self.bn3)]
        if self.downsample is not None:
            pairs.append((self.downsample[0], self.downsample[1]))
        return [(c.weight, c.shadow(), b.weight, b.bias, b.running_mean, b.running_var) for c, b in pairs]

    def bns(self):
        out = [self.bn1, self.bn2, self.bn3]
        if self.downsample is not None:
            out.append(self.downsample[1])
        return out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda and self.training and torch.is_grad_enabled():
            # whole block as one fused autograd node (ops.functional._Bottleneck)
            with torch.no_grad():
                torch._foreach_add_([b.num_batches_tracked for b in self.bns()], 1)
            return F_.bottleneck(x, self._units(), self.stride, self.bn1.eps, self.bn1.momentum)
        out = conv_bn_act(x, self.conv1, self.bn1, relu=True)
        out = conv_bn_act(out, self.conv2, self.bn2, relu=True)
        if self.downsample is not None:
            identity = conv_bn_act(x, self.downsample[0], self.downsample[1], relu=False)
        else:
            identity = F_.as_nhwc_bf16(x) if x.is_cuda else x
        return conv_bn_act(out, self.conv3, self.bn3, relu=True, residual=identity)


def stem_pool(conv: Conv2d, bn: BatchNorm2d, pool: nn.Module, x: torch.Tensor) -> torch.Tensor:
    """conv1 -> bn1 -> relu -> maxpool.  Training on a GPU with a conv output of even height and width runs the
    BN/ReLU/pool part as one fused kernel each way (``ops.functional.stem_bn_relu_pool``)."""
    if x.is_cuda:
        fuse = (bn.training and conv.is_stem and x.shape[2] % 4 == 0 and x.shape[3] % 4 == 0
                and os.environ.get("PTDB200_STEM_POOL_FUSE", "1") != "0")
        xin = F_.stem_input(x)
        if fuse:
            with torch.no_grad():
                bn.num_batches_tracked += 1
            return F_.stem_bn_relu_pool(xin, conv.weight, conv.shadow(), bn.weight, bn.bias, bn.running_mean,
                                        bn.running_var, bn.eps, bn.momentum)
        x = xin
    return pool(conv_bn_act(x, conv, bn, relu=True))


class Stem(nn.Module):
    """conv1 + bn1 + relu as one fused node (shares the parent's parameters)."""

    def __init__(self, conv: Conv2d, bn: BatchNorm2d) -> None:
        super().__init__()
        self.conv, self.bn = conv, bn

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            x = F_.stem_input(x)
        return conv_bn_act(x, self.conv, self.bn, relu=True)


class StemPool(nn.Module):
    """conv1 + bn1 + relu + maxpool (used by ``stage_modules``; shares the parent's parameters)."""

    def __init__(self, conv: Conv2d, bn: BatchNorm2d, pool: nn.Module) -> None:
        super().__init__()
        self.conv, self.bn, self.pool = conv, bn, pool

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return stem_pool(self.conv, self.bn, self.pool, x)


class ResNet(nn.Module):
    def __init__(self, block=Bottleneck, layers: Sequence[int] = (3, 4, 6, 3), num_classes: int = 1000,
                 zero_init_residual: bool = False) -> None:
        super().__init__()
        self.inplanes = 64
        self.conv1 = Conv2d(3, 64, 7, 2, 3)
        self.bn1 = BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = GlobalAvgPool()
        self.fc = Linear(512 * block.expansion, num_classes)
        if zero_init_residual:
            for m in self.modules():
                if isinstance(m, Bottleneck):
                    nn.init.constant_(m.bn3.weight, 0)

    def _make_layer(self, block, planes: int, blocks: int, stride: int = 1) -> nn.Sequential:
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(Conv2d(self.inplanes, planes * block.expansion, 1, stride),
                                       BatchNorm2d(planes * block.expansion))
        layers: List[nn.Module] = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes))
        return nn.Sequential(*layers)

    # -- pieces -----------------------------------------------------------------------------
    def stem(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            x = F_.stem_input(x)
        return conv_bn_act(x, self.conv1, self.bn1, relu=True)

    def features(self, x: torch.Tensor) -> torch.Tensor:
        x = stem_pool(self.conv1, self.bn1, self.maxpool, x)
        x = self.layer1(x)
        x = self.layer2(x)
        x = self.layer3(x)
        x = self.layer4(x)
        return x

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.avgpool(self.features(x))
        return self.fc(torch.flatten(x, 1))

    def stage_modules(self):
        """The reference's 2-stage split (notebook 03 cell5): ``seq1`` = stem..layer2, ``seq2`` =
        layer3..avgpool; ``fc`` stays separate.  The Sequentials share this model's parameters."""
        seq1 = nn.Sequential(StemPool(self.conv1, self.bn1, self.maxpool), self.layer1, self.layer2)
        seq2 = nn.Sequential(self.layer3, self.layer4, self.avgpool)
        return seq1, seq2, self.fc


def resnet50(num_classes: int = 1000, **kw) -> ResNet:
    return ResNet(Bottleneck, (3, 4, 6, 3), num_classes=num_classes, **kw)
