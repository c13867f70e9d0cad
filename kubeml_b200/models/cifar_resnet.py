"""CIFAR ResNet-20/32/44/56/110/1202 with option-A (parameter-free, zero-padded, strided) shortcuts —
the family the reference ships as `resnet32.py` (ml/experiments/kubeml/resnet32.py:44-239): 3x3 stem
with 16 channels, three stages of n BasicBlocks with 16/32/64 channels, global average pool, linear.
Stock torch modules: the compatibility path (arena-resident weights, flat optimizer, fused loss, fused
merge, cuDNN compute) and the oracle of `models.cifar_native`, which runs the same networks on the
hand-written kernels and is what the `resnet32` function uses by default."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class _ShortcutA(nn.Module):
    def __init__(self, planes: int):
        super().__init__()
        self.pad = planes // 4

    def forward(self, x):
        return F.pad(x[:, :, ::2, ::2], (0, 0, 0, 0, self.pad, self.pad), "constant", 0.0)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, in_planes: int, planes: int, stride: int = 1):
        super().__init__()
        self.conv1 = nn.Conv2d(in_planes, planes, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.shortcut = _ShortcutA(planes) if (stride != 1 or in_planes != planes) else nn.Sequential()

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return F.relu(out + self.shortcut(x))


class CifarResNet(nn.Module):
    def __init__(self, num_blocks, num_classes: int = 10):
        super().__init__()
        self.in_planes = 16
        self.conv1 = nn.Conv2d(3, 16, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(16)
        self.layer1 = self._make_layer(16, num_blocks[0], 1)
        self.layer2 = self._make_layer(32, num_blocks[1], 2)
        self.layer3 = self._make_layer(64, num_blocks[2], 2)
        self.linear = nn.Linear(64, num_classes)
        for m in self.modules():
            if isinstance(m, (nn.Linear, nn.Conv2d)):
                nn.init.kaiming_normal_(m.weight)

    def _make_layer(self, planes, n, stride):
        layers = []
        for s in [stride] + [1] * (n - 1):
            layers.append(BasicBlock(self.in_planes, planes, s))
            self.in_planes = planes
        return nn.Sequential(*layers)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.layer3(self.layer2(self.layer1(out)))
        out = F.avg_pool2d(out, out.size(3)).flatten(1)
        return self.linear(out)


def resnet20(num_classes=10): return CifarResNet([3, 3, 3], num_classes)
def resnet32(num_classes=10): return CifarResNet([5, 5, 5], num_classes)
def resnet44(num_classes=10): return CifarResNet([7, 7, 7], num_classes)
def resnet56(num_classes=10): return CifarResNet([9, 9, 9], num_classes)
def resnet110(num_classes=10): return CifarResNet([18, 18, 18], num_classes)
def resnet1202(num_classes=10): return CifarResNet([200, 200, 200], num_classes)
