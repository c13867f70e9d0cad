"""VGG-11/13/16/19 with torchvision's topology, parameter names (`features.N.weight`,
`classifier.N.bias`) and initialisation — the network the reference's VGG function instantiates
(ml/experiments/kubeml/function_vgg11.py:102, `torchvision.models.vgg.vgg11()`; BASELINE.md config 3
names VGG-16) — executed by the hand-written kernels:

  features    conv3x3 = shifted-TMA implicit GEMM on tcgen05 with bias + ReLU in the epilogue
              (the 3-channel first layer goes through im2col), 2x2 max-pool
  avgpool     AdaptiveAvgPool2d((7,7)) of the 1x1 map a 32x32 image ends on = a 49x broadcast
  classifier  GEMM + bias + ReLU epilogue -> dropout drawn in-kernel from a device step counter
              (CUDA-graph replays draw fresh masks) -> ... -> GEMM + bias
and backwards: one fused ReLU(+dropout) mask kernel per layer, bias gradient = column sum, weight
gradient on the side stream, data gradient GEMMs that read the filters MN-major (no transposes).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Union

import torch
import torch.nn as nn

from ..ops import nn as K
from .native import (Act, Buffers, NativeNet, NConv2d, NLinear, _cluster_enabled, act_dtype, conv_backward,
                     conv_forward, linear_backward, linear_forward)

CFGS: Dict[str, List[Union[int, str]]] = {
    "vgg11": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "vgg13": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "vgg16": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "vgg19": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M",
              512, 512, 512, 512, "M"],
}


class _Slot(nn.Module):
    """Parameter-less placeholder that keeps torchvision's Sequential indices (ReLU / MaxPool /
    Dropout positions) so the state_dict keys match."""

    def __init__(self, kind: str):
        super().__init__()
        self.kind = kind

    def extra_repr(self) -> str:
        return self.kind


class NativeVGG(NativeNet):
    POOL_OUT = 7                  # torchvision: nn.AdaptiveAvgPool2d((7, 7))

    def __init__(self, cfg: List[Union[int, str]], num_classes: int = 1000, dropout: float = 0.5):
        super().__init__()
        feats: List[nn.Module] = []
        cin = 3
        for v in cfg:
            if v == "M":
                feats.append(_Slot("maxpool2x2"))
            else:
                conv = NConv2d(cin, int(v), 3, 1, 1, bias=True)          # kaiming fan_out, zero bias
                feats += [conv, _Slot("relu")]
                cin = int(v)
        self.features = nn.Sequential(*feats)
        r2 = self.POOL_OUT * self.POOL_OUT
        self.classifier = nn.Sequential(
            NLinear(cin * r2, 4096), _Slot("relu"), _Slot("dropout"),
            NLinear(4096, 4096), _Slot("relu"), _Slot("dropout"),
            NLinear(4096, num_classes))
        for m in self.classifier:
            if isinstance(m, NLinear):
                nn.init.normal_(m.weight, 0, 0.01)
                nn.init.zeros_(m.bias)
        self.p = float(dropout)
        self.num_classes = num_classes
        self.feat_channels = cin
        self._saved: Optional[dict] = None
        self._counter: Optional[torch.Tensor] = None

    def _step_counter(self, device) -> torch.Tensor:
        if self._counter is None or self._counter.device != device:
            self._counter = torch.zeros(1, dtype=torch.int64, device=device)
        return self._counter

    # ---------------------------------------------------------------------------------- forward
    def _to_nhwc(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() == 4 and x.shape[-1] == 3 and x.shape[1] != 3:
            return x.contiguous().to(act_dtype())
        if x.dim() == 4 and x.shape[1] == 3:
            return x.permute(0, 2, 3, 1).contiguous().to(act_dtype())
        raise ValueError(f"unexpected input shape {tuple(x.shape)}")

    def _forward_impl(self, x: torch.Tensor, train: bool) -> torch.Tensor:
        x = self._to_nhwc(x)
        N = x.shape[0]
        B = self.buffers_for(N, train)
        tape = []                                  # (kind, module/None, input, output)
        h = x
        for i, m in enumerate(self.features):
            if isinstance(m, NConv2d):
                y = conv_forward(m, h, B, f"f{i}", bias=m.bias.data, relu=True, allow_slabs=True, fold=True).t
                tape.append(("conv", m, f"f{i}", h, y))
                h = y
            elif m.kind == "maxpool2x2":
                n_, H, W, C = h.shape
                y = B.get(f"f{i}.y", (n_, H // 2, W // 2, C))
                idx = B.get(f"f{i}.idx", (n_, H // 2, W // 2, C), torch.uint8)
                K.maxpool_fwd(h, y, idx, 2, 2, 0)
                tape.append(("pool", None, f"f{i}", h, y))
                h = y
        n_, H, W, C = h.shape
        R = self.POOL_OUT
        if H * W == 1:
            flat = B.get("flat", (N, C * R * R))
            K.replicate_fwd(h.view(N, C), flat, R * R)
        elif H == R and W == R:                    # 224x224 inputs: the pool is the identity
            flat = B.get("flat", (N, C * R * R))
            flat.copy_(h.permute(0, 3, 1, 2).reshape(N, C * R * R))
        else:
            raise ValueError(f"NativeVGG supports inputs that end on a 1x1 or {R}x{R} feature map, got {H}x{W}")
        fc1, fc2, fc3 = self.classifier[0], self.classifier[3], self.classifier[6]
        drop = train and self.p > 0
        cnt = self._step_counter(x.device) if drop else None
        a1 = linear_forward(fc1, flat, B, "c0", relu=True)
        d1 = a1
        if drop:
            d1 = B.get("c2.y", tuple(a1.shape))
            K.dropout_fwd(a1, d1, None, cnt, 1, self.p)
        a2 = linear_forward(fc2, d1, B, "c3", relu=True)
        d2 = a2
        if drop:
            d2 = B.get("c5.y", tuple(a2.shape))
            K.dropout_fwd(a2, d2, None, cnt, 2, self.p)
        logits = linear_forward(fc3, d2, B, "c6")
        if train:
            self._saved = dict(B=B, tape=tape, last=h, flat=flat, d1=d1, d2=d2, drop=drop)
        return logits

    # ---------------------------------------------------------------------------------- backward
    @staticmethod
    def _mask(dy: Act, y: torch.Tensor, g: torch.Tensor, scale: float, dbias: torch.Tensor) -> None:
        """g = scale * dy * (y > 0) and dbias = column sums of g in one pass; `dy` may be split-K slabs."""
        K.relu_bwd_cols(None if dy.slabs else dy.t, y, g, scale, dbias, dy_slabs=dy.t if dy.slabs else None)

    def _backward_impl(self, dlogits: torch.Tensor) -> None:
        S = self._saved
        B: Buffers = S["B"]
        if dlogits.dtype != act_dtype():
            dlogits = dlogits.to(act_dtype())
        self._begin_backward(B)
        fc1, fc2, fc3 = self.classifier[0], self.classifier[3], self.classifier[6]
        scale = 1.0 / (1.0 - self.p) if S["drop"] else 1.0
        # the ReLU(+dropout) mask kernel also folds split-K slabs and emits the bias gradient
        dd2 = linear_backward(fc3, S["d2"], dlogits, B, "c6", allow_slabs=True)
        g2 = B.get("c3.g", tuple(dd2.shape))
        self._mask(dd2, S["d2"], g2, scale, fc2.b_grad)
        dd1 = linear_backward(fc2, S["d1"], g2, B, "c3", bias_done=True, allow_slabs=True)
        g1 = B.get("c0.g", tuple(dd1.shape))
        self._mask(dd1, S["d1"], g1, scale, fc1.b_grad)
        dflat = linear_backward(fc1, S["flat"], g1, B, "c0", bias_done=True)
        last = S["last"]
        N, H, W, C = last.shape
        R = self.POOL_OUT
        dh_t = B.get("flat.dx", tuple(last.shape))
        if H * W == 1:
            K.replicate_bwd(dflat, dh_t.view(N, C), R * R)
        else:
            dh_t.copy_(dflat.view(N, C, R, R).permute(0, 2, 3, 1))
        dh = Act(dh_t)
        tape = S["tape"]
        for ti in range(len(tape) - 1, -1, -1):
            kind, m, tag, xin, y = tape[ti]
            if kind == "pool":
                dx = B.get(tag + ".dx", tuple(xin.shape))
                K.maxpool_bwd(dh.t, B[tag + ".idx"], dx, 2, 2, 0)
                dh = Act(dx)
            else:
                g = B.get(tag + ".g", tuple(y.shape))
                self._mask(dh, y, g, 1.0, m.b_grad)
                # split-K is welcome everywhere: the cluster reduction delivers a dense bf16 gradient
                # (fp32 slabs, the fallback, only when the consumer is another mask kernel)
                dh = conv_backward(m, xin, g, B, tag, need_dx=ti > 0,
                                   allow_slabs=ti > 0 and (_cluster_enabled() or tape[ti - 1][0] == "conv"))
        if S["drop"]:
            K.bump_counter(self._counter)
        self._end_backward(B)


def _vgg(name: str, num_classes: int, dropout: float) -> NativeVGG:
    return NativeVGG(CFGS[name], num_classes, dropout)


def vgg11(num_classes: int = 1000, dropout: float = 0.5) -> NativeVGG:
    return _vgg("vgg11", num_classes, dropout)


def vgg13(num_classes: int = 1000, dropout: float = 0.5) -> NativeVGG:
    return _vgg("vgg13", num_classes, dropout)


def vgg16(num_classes: int = 1000, dropout: float = 0.5) -> NativeVGG:
    return _vgg("vgg16", num_classes, dropout)


def vgg19(num_classes: int = 1000, dropout: float = 0.5) -> NativeVGG:
    return _vgg("vgg19", num_classes, dropout)
