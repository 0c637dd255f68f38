from .layers import BatchNorm2d, Conv2d, GlobalAvgPool, Linear, MaxPool2d, conv_bn_act
from .resnet import Bottleneck, ResNet, resnet50
from .toy import MLP, ToyLinear

__all__ = ["BatchNorm2d", "Conv2d", "GlobalAvgPool", "Linear", "MaxPool2d", "conv_bn_act", "Bottleneck", "ResNet",
           "resnet50", "MLP", "ToyLinear"]
