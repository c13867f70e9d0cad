"""CIFAR ResNet-20/32/44/56/110 (option-A shortcuts) on the hand-written kernels — the family the
reference ships as `resnet32.py` (ml/experiments/kubeml/resnet32.py:44-239) and trains in its
ResNet-32 / CIFAR-10 experiments.  Same parameter names and shapes as `models.cifar_resnet`
(`conv1`, `bn1`, `layer{1,2,3}.{i}.{conv1,bn1,conv2,bn2}`, `linear`).

The 16- and 32-channel stages are below the 64-channel granularity of the shifted-TMA implicit GEMM:
their convolutions run as im2col + tcgen05 GEMM (cluster split-K, BatchNorm statistics in the
epilogue), the 64-channel stage takes the direct path.  The parameter-free shortcut of a down-sampling
block, `pad(x[:, :, ::2, ::2], planes // 4 channels on both sides)`, is a strided copy into a zeroed
buffer; its gradient is the same copy backwards, handed to the convolution's col2im as the addend.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn as nn

from ..ops import nn as K
from .native import (Act, Buffers, NativeNet, NBatchNorm2d, NConv2d, NLinear, act_dtype, bn_backward, bn_forward,
                     conv_backward, conv_bn_forward, conv_forward, linear_backward, linear_forward)


class NCifarBlock(nn.Module):
    expansion = 1

    def __init__(self, cin: int, planes: int, stride: int):
        super().__init__()
        self.conv1 = NConv2d(cin, planes, 3, stride, 1)
        self.bn1 = NBatchNorm2d(planes)
        self.conv2 = NConv2d(planes, planes, 3, 1, 1)
        self.bn2 = NBatchNorm2d(planes)
        self.shortcut = nn.Sequential()              # option A: no parameters (key-compatible placeholder)
        self.pad = planes // 4 if (stride != 1 or cin != planes) else 0
        self.stride = stride


class NativeCifarResNet(NativeNet):
    def __init__(self, num_blocks: List[int], num_classes: int = 10):
        super().__init__()
        self.conv1 = NConv2d(3, 16, 3, 1, 1)
        self.bn1 = NBatchNorm2d(16)
        cin = 16
        for i, (planes, n, stride) in enumerate(zip((16, 32, 64), num_blocks, (1, 2, 2)), start=1):
            blocks = []
            for s in [stride] + [1] * (n - 1):
                blocks.append(NCifarBlock(cin, planes, s))
                cin = planes
            setattr(self, f"layer{i}", nn.Sequential(*blocks))
        self.linear = NLinear(64, num_classes)
        for m in self.modules():                      # the reference applies kaiming_normal_ to conv and linear
            if isinstance(m, (NConv2d, NLinear)):
                nn.init.kaiming_normal_(m.weight)
        self.num_classes = num_classes
        self._blocks = [(f"layer{i}.{j}", blk) for i in range(1, 4) for j, blk in enumerate(getattr(self, f"layer{i}"))]
        self._saved: Optional[dict] = None

    # ---------------------------------------------------------------------------------- forward
    @staticmethod
    def _to_nhwc(x: torch.Tensor) -> torch.Tensor:
        if x.dim() == 4 and x.shape[-1] == 3 and x.shape[1] != 3:
            return x.contiguous().to(act_dtype())
        if x.dim() == 4 and x.shape[1] == 3:
            return x.permute(0, 2, 3, 1).contiguous().to(act_dtype())
        raise ValueError(f"unexpected input shape {tuple(x.shape)}")

    def _forward_impl(self, x: torch.Tensor, train: bool) -> torch.Tensor:
        x = self._to_nhwc(x)
        N = x.shape[0]
        B = self.buffers_for(N, train)
        h, c1s = conv_bn_forward(self.conv1, self.bn1, x, B, "stem", "stem.bn", train, relu=True)
        saved = []
        for tag, blk in self._blocks:
            idn = h
            if blk.pad:
                n_, H, W, C = h.shape
                idn = B.get(tag + ".sc", (n_, H // 2, W // 2, C + 2 * blk.pad), zero=True)
                idn[..., blk.pad:blk.pad + C].copy_(h[:, ::2, ::2, :])
            # convolution + BatchNorm (+ shortcut + ReLU) in one launch where the GEMM's grid fits the GPU at once
            a, t1s = conv_bn_forward(blk.conv1, blk.bn1, h, B, tag + ".c1", tag + ".bn1", train, relu=True)
            out, t2s = conv_bn_forward(blk.conv2, blk.bn2, a, B, tag + ".c2", tag + ".bn2", train, relu=True, res=idn)
            saved.append((h, t1s, a, t2s, out))
            h = out
        n_, H, W, C = h.shape
        pooled = B.get("avg.y", (n_, C))
        K.avgpool_fwd(h, pooled)
        logits = linear_forward(self.linear, pooled, B, "fc")
        if train:
            self._saved = dict(B=B, x=x, c1=c1s, a1=saved[0][0], blocks=saved, last=h, pooled=pooled)
        return logits

    # ---------------------------------------------------------------------------------- backward
    def _backward_impl(self, dlogits: torch.Tensor) -> None:
        S = self._saved
        B: Buffers = S["B"]
        if dlogits.dtype != act_dtype():
            dlogits = dlogits.to(act_dtype())
        self._begin_backward(B)
        dpooled = linear_backward(self.linear, S["pooled"], dlogits, B, "fc")
        t = B.get("avg.dx", tuple(S["last"].shape))
        K.avgpool_bwd(dpooled, t)
        dh = Act(t)
        for (tag, blk), (h, t1, a, t2, out) in zip(reversed(self._blocks), reversed(S["blocks"])):
            dt2, g = bn_backward(blk.bn2, dh, out, t2, B, tag + ".bn2", want_dres=True)
            da = conv_backward(blk.conv2, a, dt2, B, tag + ".c2", allow_slabs=True)
            dt1, _ = bn_backward(blk.bn1, da, a, t1, B, tag + ".bn1")
            if blk.pad:                               # shortcut gradient: the strided copy backwards
                add = B.get(tag + ".scg", tuple(h.shape), zero=True)
                C = h.shape[-1]
                add[:, ::2, ::2, :].copy_(g[..., blk.pad:blk.pad + C])
            else:
                add = g
            dh = conv_backward(blk.conv1, h, dt1, B, tag + ".c1", addend=add, allow_slabs=True)
        dc1, _ = bn_backward(self.bn1, dh, S["a1"], S["c1"], B, "stem.bn")
        conv_backward(self.conv1, S["x"], dc1, B, "stem", need_dx=False)
        self._end_backward(B)


def _make(n: int):
    def ctor(num_classes: int = 10) -> NativeCifarResNet:
        return NativeCifarResNet([n, n, n], num_classes)
    return ctor


resnet20, resnet32, resnet44, resnet56, resnet110 = _make(3), _make(5), _make(7), _make(9), _make(18)
resnet1202 = _make(200)       # the reference's resnet32.py:233-234 lists it too (19.4 M parameters, 1202 layers)
