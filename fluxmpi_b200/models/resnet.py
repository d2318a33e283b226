"""ResNet-50 (v1.5 bottleneck, 25,557,032 parameters) — the model of the headline benchmark
(BASELINE.json: "ResNet-50 synthetic 224x224 ImageNet-shape data-parallel"). The reference
ships no model code at all (model math lives in Lux/Flux); its README points to an
external ImageNet example (/root/reference/README.md:78).

``fused_bn=True`` replaces every BatchNorm2d (+ the ReLU and the residual add behind it) by
:class:`fluxmpi_b200.ops.batchnorm.FusedBatchNormAct` — hand-written channels-last kernels —
because on B200 those memory-bound ops, not the convolutions, are ~75 % of the step
(profiles/README.md §7). Parameter names, shapes and state-dict keys are identical either way.
Convolutions stay in PyTorch/cuDNN.
"""
from __future__ import annotations

import torch
from torch import nn

from ..ops.batchnorm import FusedBatchNormAct
from ..ops.pool import FusedMaxPool2d


def _bn(planes: int, relu: bool, fused: bool) -> nn.Module:
    return FusedBatchNormAct(planes, relu=relu) if fused else nn.BatchNorm2d(planes)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: nn.Module | None = None,
                 fused_bn: bool = False):
        super().__init__()
        self.fused_bn = fused_bn
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = _bn(planes, True, fused_bn)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride=stride, padding=1, bias=False)
        self.bn2 = _bn(planes, True, fused_bn)
        self.conv3 = nn.Conv2d(planes, planes * self.expansion, 1, bias=False)
        self.bn3 = _bn(planes * self.expansion, True, fused_bn)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample

    def forward(self, x):
        identity = x if self.downsample is None else self.downsample(x)
        if self.fused_bn:
            out = self.bn1(self.conv1(x))                       # bn + relu in one pass
            out = self.bn2(self.conv2(out))
            return self.bn3(self.conv3(out), identity)          # bn + add + relu in one pass
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        return self.relu(out + identity)


class ResNet(nn.Module):
    def __init__(self, layers=(3, 4, 6, 3), num_classes: int = 1000, width: int = 64, fused_bn: bool = False):
        super().__init__()
        self.fused_bn = fused_bn
        self.inplanes = width
        self.conv1 = nn.Conv2d(3, width, 7, stride=2, padding=3, bias=False)
        self.bn1 = _bn(width, True, fused_bn)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = FusedMaxPool2d(3, 2, 1) if fused_bn else nn.MaxPool2d(3, stride=2, padding=1)
        self.layer1 = self._make_layer(width, layers[0])
        self.layer2 = self._make_layer(width * 2, layers[1], stride=2)
        self.layer3 = self._make_layer(width * 4, layers[2], stride=2)
        self.layer4 = self._make_layer(width * 8, layers[3], stride=2)
        self.avgpool = nn.AdaptiveAvgPool2d(1)
        self.fc = nn.Linear(width * 8 * Bottleneck.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
        for m in self.modules():
            if isinstance(m, Bottleneck):
                nn.init.zeros_(m.bn3.weight)

    def _make_layer(self, planes: int, blocks: int, stride: int = 1):
        downsample = None
        if stride != 1 or self.inplanes != planes * Bottleneck.expansion:
            downsample = nn.Sequential(
                nn.Conv2d(self.inplanes, planes * Bottleneck.expansion, 1, stride=stride, bias=False),
                _bn(planes * Bottleneck.expansion, False, self.fused_bn))
        layers = [Bottleneck(self.inplanes, planes, stride, downsample, self.fused_bn)]
        self.inplanes = planes * Bottleneck.expansion
        layers += [Bottleneck(self.inplanes, planes, fused_bn=self.fused_bn)
                   for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    def forward(self, x):
        x = self.bn1(self.conv1(x))
        if not self.fused_bn:
            x = self.relu(x)
        x = self.maxpool(x)
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return self.fc(torch.flatten(self.avgpool(x), 1))


def resnet50(num_classes: int = 1000, fused_bn: bool = False) -> ResNet:
    return ResNet((3, 4, 6, 3), num_classes, fused_bn=fused_bn)


def resnet_tiny(num_classes: int = 10, fused_bn: bool = False) -> ResNet:
    """Same topology, 1 block per stage, width 16 — for smoke tests."""
    return ResNet((1, 1, 1, 1), num_classes, width=16, fused_bn=fused_bn)
