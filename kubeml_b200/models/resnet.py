"""ResNet-18/34 (BasicBlock) with torchvision's topology, parameter names and initialisation —
the network `torchvision.models.resnet.resnet34()` that the reference's function instantiates
(ml/experiments/kubeml/function_resnet34.py:101), executed by the hand-written kernels.

Per step on a [N,32,32,3] batch the program is, in kernels:
  stem      im2col -> tcgen05 GEMM -> BN+ReLU (coop) -> max-pool
  block     conv (shifted-TMA implicit GEMM, or im2col+GEMM when stride 2) -> BN+ReLU ->
            conv -> BN + residual + ReLU        [+ 1x1/2 conv -> BN on the identity when downsampling]
  head      (avg-pool) -> GEMM+bias
and the mirror image backwards (BN backward produces the masked residual gradient, the first
convolution's data gradient adds it in its epilogue, weight gradients land in the arena).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn as nn

from .. import ops
from ..ops import nn as K
from .native import (Buffers, NativeNet, NBatchNorm2d, NConv2d, NLinear, bn_backward, bn_forward,
                     conv_backward, conv_bn_forward, conv_forward, linear_backward, linear_forward)


class NBasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin: int, cout: int, stride: int):
        super().__init__()
        self.conv1 = NConv2d(cin, cout, 3, stride, 1)
        self.bn1 = NBatchNorm2d(cout)
        self.conv2 = NConv2d(cout, cout, 3, 1, 1)
        self.bn2 = NBatchNorm2d(cout)
        self.downsample = None
        if stride != 1 or cin != cout:
            self.downsample = nn.Sequential(NConv2d(cin, cout, 1, stride, 0), NBatchNorm2d(cout))


class NativeResNet(NativeNet):
    def __init__(self, layers: List[int], num_classes: int = 1000):
        super().__init__()
        self.conv1 = NConv2d(3, 64, 7, 2, 3)
        self.bn1 = NBatchNorm2d(64)
        cin = 64
        for i, (n, c) in enumerate(zip(layers, (64, 128, 256, 512)), start=1):
            blocks = []
            for b in range(n):
                blocks.append(NBasicBlock(cin, c, 2 if (b == 0 and i > 1) else 1))
                cin = c
            setattr(self, f"layer{i}", nn.Sequential(*blocks))
        self.fc = NLinear(512, num_classes)
        self.num_classes = num_classes
        self._blocks = [(f"layer{i}.{j}", blk) for i in range(1, 5)
                        for j, blk in enumerate(getattr(self, f"layer{i}"))]
        self._saved: Optional[dict] = None
        import os
        self._stamps_on = os.environ.get("KUBEML_STAMPS") == "1"
        self.stamps: Optional[torch.Tensor] = None

    # wgrad reduces over N*Ho*Wo pixels: shallow-and-wide early layers need K-splits to fill the GPU
    @staticmethod
    def _wgrad_splits(pixels: int, cout: int, cin_cols: int) -> int:
        tiles = ((cout + 127) // 128) * max(1, cin_cols // 64)
        chunks = max(1, pixels // 64)
        want = max(1, min(chunks // 4, 148 // max(tiles, 1)))
        return want

    def _after_bound(self, arena) -> None:
        # gradients produced with split-K are accumulated with red.add and must start from zero:
        # remember the flat range that covers them (the early, small layers)
        self._zero_names = []

    # ---------------------------------------------------------------------------------- forward
    def _to_nhwc_bf16(self, x: torch.Tensor) -> torch.Tensor:
        from .native import act_dtype
        if x.dim() == 4 and x.shape[-1] == 3 and x.shape[1] != 3:
            return x.contiguous().to(act_dtype())
        if x.dim() == 4 and x.shape[1] == 3:          # NCHW float (reference-style pipeline)
            return x.permute(0, 2, 3, 1).contiguous().to(act_dtype())
        raise ValueError(f"unexpected input shape {tuple(x.shape)}")

    # Optional phase timeline (KUBEML_STAMPS=1, bench/step_timeline.py): the GPU timer is written into
    # `self.stamps[i]` at stage boundaries of the forward and backward programs.
    STAMP_NAMES = ["start", "stem", "layer1", "layer2", "layer3", "layer4", "head_fwd", "bwd_head", "bwd_layer4",
                   "bwd_layer3", "bwd_layer2", "bwd_layer1", "bwd_stem_end"]

    def _stamp(self, name: str) -> None:
        if not self._stamps_on:
            return
        if self.stamps is None or self.stamps.device != self._arena.weights.device:
            self.stamps = torch.zeros(32, dtype=torch.int64, device=self._arena.weights.device)
        K.stamp(self.stamps, self.STAMP_NAMES.index(name))

    def _forward_impl(self, x: torch.Tensor, train: bool) -> torch.Tensor:
        x = self._to_nhwc_bf16(x)
        N = x.shape[0]
        B = self.buffers_for(N, train)
        self._stamp("start")
        S = {"x": x, "blocks": []}
        a1, c1s = conv_bn_forward(self.conv1, self.bn1, x, B, "stem", "stem.bn", train, relu=True)
        _, H1, W1, _ = a1.shape
        Hp, Wp = (H1 + 2 - 3) // 2 + 1, (W1 + 2 - 3) // 2 + 1
        p1 = B.get("pool.y", (N, Hp, Wp, 64))
        idx = B.get("pool.idx", (N, Hp, Wp, 64), torch.uint8)
        K.maxpool_fwd(a1, p1, idx, 3, 2, 1)
        S.update(c1=c1s, a1=a1, p1=p1)
        h = p1
        self._stamp("stem")
        for tag, blk in self._blocks:
            if tag.endswith(".0") and not tag.startswith("layer1"):
                self._stamp("layer" + str(int(tag[5]) - 1))
            idn, tds = h, None
            if blk.downsample is not None:       # the projection shortcut runs next to the main branch
                with B.on_side():
                    td = conv_forward(blk.downsample[0], h, B, tag + ".cd", allow_slabs=True, want_stats=train)
                    idn, tds = bn_forward(blk.downsample[1], td, B, tag + ".bnd", train, relu=False)
            a, t1s = conv_bn_forward(blk.conv1, blk.bn1, h, B, tag + ".c1", tag + ".bn1", train, relu=True)
            if blk.downsample is not None:
                B.join_side()
            out, t2s = conv_bn_forward(blk.conv2, blk.bn2, a, B, tag + ".c2", tag + ".bn2", train, relu=True, res=idn)
            S["blocks"].append((h, t1s, a, t2s, tds, out))
            h = out
        self._stamp("layer4")
        Nn, Hh, Ww, C = h.shape
        if Hh * Ww == 1:
            pooled = h.view(Nn, C)
        else:
            pooled = B.get("avg.y", (Nn, C))
            K.avgpool_fwd(h, pooled)
        logits = linear_forward(self.fc, pooled, B, "fc")
        self._stamp("head_fwd")
        S.update(last=h, pooled=pooled, B=B)
        if train:
            self._saved = S
        return logits

    def _mark_tail_ready(self, B: Buffers) -> None:
        """Called in the backward pass once every gradient of layer4 and fc has been issued: publishes
        (flat offset of layer4's first parameter, event on the main stream, event on the weight-gradient stream)
        for the flat optimizer, which then updates that tail of the arena while the rest of the backward pass
        still runs (`kubeml_b200.optim.SGD.step`)."""
        a = self._arena
        if a is None or a.weights.device.type != "cuda":
            return
        lo = a.layout.entries["layer4.0.conv1.weight"].offset
        if lo % 64:
            return
        dev = a.weights.device
        ev_main, ev_side = torch.cuda.Event(), torch.cuda.Event()
        ev_main.record(torch.cuda.current_stream(dev))
        side = getattr(B, "_side", None)
        ev_side.record(side if (side is not None and getattr(B, "_side_used", False)) else torch.cuda.current_stream(dev))
        a.early_ready = (lo, ev_main, ev_side)

    # ---------------------------------------------------------------------------------- backward
    def _backward_impl(self, dlogits: torch.Tensor) -> None:
        from .native import Act, act_dtype
        S = self._saved
        B: Buffers = S["B"]
        if dlogits.dtype != act_dtype():
            dlogits = dlogits.to(act_dtype())
        self._begin_backward(B)
        dpooled = linear_backward(self.fc, S["pooled"], dlogits, B, "fc")
        last = S["last"]
        Nn, Hh, Ww, C = last.shape
        if Hh * Ww == 1:
            dh = Act(dpooled.view(Nn, 1, 1, C))
        else:
            t = B.get("avg.dx", tuple(last.shape))
            K.avgpool_bwd(dpooled, t)
            dh = Act(t)
        nblk = len(self._blocks)
        from .native import sm_count
        wb = sm_count() // 2                 # weight gradients have slack here: leave half of the SMs to the chain
        self._stamp("bwd_head")
        for bi, ((tag, blk), (h, t1, a, t2, td, out)) in enumerate(zip(reversed(self._blocks),
                                                                         reversed(S["blocks"]))):
            if bi > 0 and tag.endswith("." + str(len(getattr(self, tag[:6])) - 1)):
                self._stamp("bwd_layer" + str(int(tag[5]) + 1))
            if tag == "layer3." + str(len(self.layer3) - 1):
                self._mark_tail_ready(B)          # layer4 and fc are done: their optimizer update may start
            first_block = bi == nblk - 1          # its input gradient feeds the max-pool backward (dense)
            # out = relu(bn2(t2) + idn)
            dt2, g = bn_backward(blk.bn2, dh, out, t2, B, tag + ".bn2", want_dres=True)
            # the data-gradient GEMMs mask and pre-reduce for the BatchNorm backward that follows them
            da = conv_backward(blk.conv2, a, dt2, B, tag + ".c2", allow_slabs=True,
                               bn_next=(blk.bn1, tag + ".bn1", a, t1), wgrad_budget=wb)
            dt1, _ = bn_backward(blk.bn1, da, a, t1, B, tag + ".bn1")
            if blk.downsample is None:
                nxt = None
                if not first_block:              # consumer: bn2 of the block before (its output is `h`)
                    ptag, pblk = self._blocks[nblk - 2 - bi]
                    nxt = (pblk.bn2, ptag + ".bn2", h, S["blocks"][nblk - 2 - bi][3], True)
                dh = conv_backward(blk.conv1, h, dt1, B, tag + ".c1", addend=g, allow_slabs=not first_block,
                                   bn_next=nxt, wgrad_budget=wb)
            else:
                dtd, _ = bn_backward(blk.downsample[1], Act(g), None, td, B, tag + ".bnd")
                did = conv_backward(blk.downsample[0], h, dtd, B, tag + ".cd", wgrad_budget=wb)
                dh = conv_backward(blk.conv1, h, dt1, B, tag + ".c1", addend=did.t, wgrad_budget=wb)
        # stem
        self._stamp("bwd_layer1")
        da1 = B.get("pool.dx", tuple(S["a1"].shape))
        K.maxpool_bwd(dh.dense(B, "pool.dy"), B["pool.idx"], da1, 3, 2, 1)
        dc1, _ = bn_backward(self.bn1, Act(da1), S["a1"], S["c1"], B, "stem.bn")
        conv_backward(self.conv1, S["x"], dc1, B, "stem", need_dx=False)
        self._end_backward(B)
        self._stamp("bwd_stem_end")


def resnet34(num_classes: int = 1000) -> NativeResNet:
    return NativeResNet([3, 4, 6, 3], num_classes)


def resnet18(num_classes: int = 1000) -> NativeResNet:
    return NativeResNet([2, 2, 2, 2], num_classes)
