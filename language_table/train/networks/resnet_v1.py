"""ResNet V1 (reference ``language_table/train/networks/resnet_v1.py``, Flax) in PyTorch, NCHW.

Same architecture and the same parameter counts as the reference's test asserts (``resnet_v1_test.py:27-41``):
7x7 s2 stem without bias, BatchNorm (momentum 0.9 in Flax terms = 0.1 in PyTorch terms), 3x3 s2 SAME max-pool, stages of
basic or bottleneck blocks with a projection shortcut where the shape changes, zero-initialised last BN scale of the basic
block and zero-initialised classifier.  ``MultiscaleResNet`` returns the feature pyramid the LAVA encoder consumes.
"""
from functools import partial
from typing import List, Sequence, Type

import torch
import torch.nn as nn
import torch.nn.functional as F


def Conv1x1(cin: int, cout: int, stride: int = 1) -> nn.Conv2d:
    """Bias-free 1×1 convolution (the reference's ``Conv1x1`` partial, ``resnet_v1.py:33``)."""
    return nn.Conv2d(cin, cout, 1, stride, 0, bias=False)


def Conv3x3(cin: int, cout: int, stride: int = 1) -> nn.Conv2d:
    """Bias-free 3×3 SAME convolution (``resnet_v1.py:34``)."""
    return nn.Conv2d(cin, cout, 3, stride, 1, bias=False)


def _bn(c: int, zero_scale: bool = False) -> nn.BatchNorm2d:
    bn = nn.BatchNorm2d(c, momentum=0.1, eps=1e-5)
    if zero_scale:
        nn.init.zeros_(bn.weight)
    return bn


class ResNetBlock(nn.Module):
    expansion = 1

    def __init__(self, cin: int, filters: int, stride: int = 1):
        super().__init__()
        self.conv1 = Conv3x3(cin, filters, stride)
        self.bn1 = _bn(filters)
        self.conv2 = Conv3x3(filters, filters)
        self.bn2 = _bn(filters, zero_scale=True)
        self.proj_conv = self.proj_bn = None
        if stride != 1 or cin != filters:
            self.proj_conv = Conv1x1(cin, filters, stride)
            self.proj_bn = _bn(filters)

    def forward(self, x):
        r = x
        y = F.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        if self.proj_conv is not None:
            r = self.proj_bn(self.proj_conv(r))
        return F.relu(r + y)


class BottleneckResNetBlock(nn.Module):
    expansion = 4

    def __init__(self, cin: int, filters: int, stride: int = 1):
        super().__init__()
        self.conv1 = Conv1x1(cin, filters)
        self.bn1 = _bn(filters)
        self.conv2 = Conv3x3(filters, filters, stride)
        self.bn2 = _bn(filters)
        self.conv3 = Conv1x1(filters, 4 * filters)
        self.bn3 = _bn(4 * filters)
        self.proj_conv = self.proj_bn = None
        if stride != 1 or cin != 4 * filters:
            self.proj_conv = Conv1x1(cin, 4 * filters, stride)
            self.proj_bn = _bn(4 * filters)

    def forward(self, x):
        r = x
        y = F.relu(self.bn1(self.conv1(x)))
        y = F.relu(self.bn2(self.conv2(y)))
        y = self.bn3(self.conv3(y))
        if self.proj_conv is not None:
            r = self.proj_bn(self.proj_conv(r))
        return F.relu(r + y)


class ResNetStage(nn.Sequential):
    """``stage_size`` blocks of ``filters``; only the first one strides (reference ``resnet_v1.py:96-114``).  PyTorch
    modules need their input width up front, hence ``cin``; ``out_channels`` is what the next stage gets."""

    def __init__(self, cin: int, stage_size: int, filters: int, block_cls: Type[nn.Module], first_block_strides: int = 1):
        blocks = []
        for j in range(stage_size):
            blocks.append(block_cls(cin, filters, stride=first_block_strides if j == 0 else 1))
            cin = filters * block_cls.expansion
        super().__init__(*blocks)
        self.out_channels = cin


def _stages(block_cls: Type[nn.Module], stage_sizes: Sequence[int], width: int):
    stages, cin = nn.ModuleList(), width
    for i, n in enumerate(stage_sizes):
        stages.append(ResNetStage(cin, n, width * 2 ** i, block_cls, first_block_strides=1 if i == 0 else 2))
        cin = stages[-1].out_channels
    return stages, cin


class ResNet(nn.Module):
    def __init__(self, num_classes: int, block_cls: Type[nn.Module], stage_sizes: Sequence[int], width_factor: int = 1):
        super().__init__()
        width = 64 * width_factor
        self.init_conv = nn.Conv2d(3, width, 7, 2, 3, bias=False)
        self.init_bn = _bn(width)
        self.stages, cout = _stages(block_cls, stage_sizes, width)
        self.head = nn.Linear(cout, num_classes)
        nn.init.zeros_(self.head.weight)
        nn.init.zeros_(self.head.bias)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.init_bn(self.init_conv(x))), 3, 2, 1)
        for s in self.stages:
            x = s(x)
        return self.head(x.mean(dim=(2, 3)))


ResNet18 = partial(ResNet, stage_sizes=(2, 2, 2, 2), block_cls=ResNetBlock)
ResNet34 = partial(ResNet, stage_sizes=(3, 4, 6, 3), block_cls=ResNetBlock)
ResNet50 = partial(ResNet, stage_sizes=(3, 4, 6, 3), block_cls=BottleneckResNetBlock)
ResNet101 = partial(ResNet, stage_sizes=(3, 4, 23, 3), block_cls=BottleneckResNetBlock)
ResNet152 = partial(ResNet, stage_sizes=(3, 8, 36, 3), block_cls=BottleneckResNetBlock)
ResNet200 = partial(ResNet, stage_sizes=(3, 24, 36, 3), block_cls=BottleneckResNetBlock)


class MultiscaleResNet(nn.Module):
    """Returns [stem conv output, pooled stem, stage 1, stage 2, ...] (reference ``:201-259``)."""

    def __init__(self, block_cls: Type[nn.Module] = BottleneckResNetBlock, stage_sizes: Sequence[int] = (3, 4), width_factor: int = 1):
        super().__init__()
        width = 64 * width_factor
        self.init_conv = nn.Conv2d(3, width, 7, 2, 3, bias=False)
        self.init_bn = _bn(width)
        self.stages, self.out_channels = _stages(block_cls, stage_sizes, width)

    def forward(self, x) -> List[torch.Tensor]:
        outs = []
        x = self.init_conv(x)
        outs.append(x)
        x = F.max_pool2d(F.relu(self.init_bn(x)), 3, 2, 1)
        outs.append(x)
        for s in self.stages:
            x = s(x)
            outs.append(x)
        return outs
