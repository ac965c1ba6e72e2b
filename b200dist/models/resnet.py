"""ResNet-18/34/50/101/152 with torchvision-compatible state-dict keys
(``conv1``, ``bn1``, ``layer{1..4}.{i}.conv{1,2,3}``, ``.bn{1,2,3}``, ``.downsample.{0,1}``, ``fc``), so
``torchvision.models.resnet50().state_dict()`` loads directly (reference: /root/reference/scripts/main.py:249
``torchvision.models.<arch>(pretrained=False)``; CIFAR stem variant of
scripts/02_fully_sharded_fsdp/resnet_fsdp_training.py:186-190 via ``cifar_stem=True``).

Execution: channels-last activations; 1x1 convs are direct tcgen05 GEMMs, kxk convs im2col + GEMM;
BatchNorm + ReLU (and the bottleneck's residual add) are one fused pass.
"""
from __future__ import annotations

from typing import List, Type, Union

import torch
from torch import nn

from .layers import BatchNorm2d, Conv2d, GlobalAvgPool, Linear, MaxPool2d


def conv3x3(i, o, stride=1):
    return Conv2d(i, o, 3, stride, 1, bias=False)


def conv1x1(i, o, stride=1):
    return Conv2d(i, o, 1, stride, 0, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1, self.bn1 = conv3x3(inplanes, planes, stride), BatchNorm2d(planes)
        self.conv2, self.bn2 = conv3x3(planes, planes), BatchNorm2d(planes)
        self.downsample = downsample

    def forward(self, x):
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x), relu=True)
        return self.bn2(self.conv2(out), residual=identity, relu=True)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1, self.bn1 = conv1x1(inplanes, planes), BatchNorm2d(planes)
        self.conv2, self.bn2 = conv3x3(planes, planes, stride), BatchNorm2d(planes)
        self.conv3, self.bn3 = conv1x1(planes, planes * 4), BatchNorm2d(planes * 4)
        self.downsample = downsample

    def forward(self, x):
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x), relu=True)
        out = self.bn2(self.conv2(out), relu=True)
        return self.bn3(self.conv3(out), residual=identity, relu=True)


class _Downsample(nn.Sequential):
    def forward(self, x):
        return self[1](self[0](x))


class ResNet(nn.Module):
    def __init__(self, block: Type[Union[BasicBlock, Bottleneck]], layers: List[int], num_classes: int = 1000,
                 cifar_stem: bool = False):
        super().__init__()
        self.inplanes = 64
        self.cifar_stem = cifar_stem
        self.conv1 = Conv2d(3, 64, 3, 1, 1, bias=False) if cifar_stem else Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = BatchNorm2d(64)
        self.maxpool = nn.Identity() if cifar_stem else MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], 2)
        self.layer3 = self._make_layer(block, 256, layers[2], 2)
        self.layer4 = self._make_layer(block, 512, layers[3], 2)
        self.avgpool = GlobalAvgPool()
        self.fc = Linear(512 * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def _make_layer(self, block, planes, blocks, stride=1):
        down = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            down = _Downsample(conv1x1(self.inplanes, planes * block.expansion, stride), BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, down)]
        self.inplanes = planes * block.expansion
        layers += [block(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    def forward(self, x):
        if x.is_cuda:
            x = x.contiguous(memory_format=torch.channels_last)
        x = self.maxpool(self.bn1(self.conv1(x), relu=True))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return self.fc(self.avgpool(x))


_CFG = {"resnet18": (BasicBlock, [2, 2, 2, 2]), "resnet34": (BasicBlock, [3, 4, 6, 3]),
        "resnet50": (Bottleneck, [3, 4, 6, 3]), "resnet101": (Bottleneck, [3, 4, 23, 3]),
        "resnet152": (Bottleneck, [3, 8, 36, 3])}


def build_resnet(arch: str = "resnet18", num_classes: int = 1000, cifar_stem: bool = False) -> ResNet:
    block, layers = _CFG[arch]
    return ResNet(block, layers, num_classes, cifar_stem)


def resnet18(**kw): return build_resnet("resnet18", **kw)
def resnet50(**kw): return build_resnet("resnet50", **kw)
def resnet101(**kw): return build_resnet("resnet101", **kw)
def resnet152(**kw): return build_resnet("resnet152", **kw)
