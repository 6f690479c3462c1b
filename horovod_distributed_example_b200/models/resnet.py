"""ResNet-50 (He et al. 2015) in plain ``torch.nn`` for the large-bucket stress test
(BASELINE.json config 5: "ResNet-50 synthetic ImageNet-shape through the same DistributedOptimizer").

The reference has no ResNet; this model exists to push ~25.6 M parameters / ~102 MB of gradients in
~161 tensors through the arena, the static bucket plan and the fused peer-memory allreduce+update, the
way Horovod's own synthetic benchmark does.  Compute here is stock torch/cuDNN by design - the
north star's "no cuDNN" clause lists the MNIST ops; the point of this model is the optimizer path.
"""
from __future__ import annotations

import torch
import torch.nn as nn


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, width, stride=1, downsample=None):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, width, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(width)
        self.conv2 = nn.Conv2d(width, width, 3, stride, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(width)
        self.conv3 = nn.Conv2d(width, width * 4, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(width * 4)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample

    def forward(self, x):
        idt = x if self.downsample is None else self.downsample(x)
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        return self.relu(out + idt)


class ResNet(nn.Module):
    def __init__(self, layers=(3, 4, 6, 3), num_classes=1000, width=64):
        super().__init__()
        self.inplanes = width
        self.conv1 = nn.Conv2d(3, width, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(width)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(3, 2, 1)
        self.layer1 = self._make(width, layers[0], 1)
        self.layer2 = self._make(width * 2, layers[1], 2)
        self.layer3 = self._make(width * 4, layers[2], 2)
        self.layer4 = self._make(width * 8, layers[3], 2)
        self.avgpool = nn.AdaptiveAvgPool2d(1)
        self.fc = nn.Linear(width * 8 * 4, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def _make(self, width, blocks, stride):
        down = None
        if stride != 1 or self.inplanes != width * 4:
            down = nn.Sequential(nn.Conv2d(self.inplanes, width * 4, 1, stride, bias=False), nn.BatchNorm2d(width * 4))
        mods = [Bottleneck(self.inplanes, width, stride, down)]
        self.inplanes = width * 4
        mods += [Bottleneck(self.inplanes, width) for _ in range(1, blocks)]
        return nn.Sequential(*mods)

    def forward(self, x):
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return self.fc(torch.flatten(self.avgpool(x), 1))


def resnet50(num_classes=1000) -> ResNet:
    return ResNet((3, 4, 6, 3), num_classes)


def resnet_tiny(num_classes=10) -> ResNet:
    """Same topology, 1 block per stage and width 8: for CPU tests."""
    return ResNet((1, 1, 1, 1), num_classes, width=8)
