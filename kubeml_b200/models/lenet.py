"""The reference's LeNet variant (ml/experiments/kubeml/function_lenet.py:21-50: conv 1->6 k5, pool,
conv 6->16 k5, pool, fc 256->120->84->10 with a ReLU after EVERY layer, the last one included) on the
hand-written kernels.

The shapes are far below the tensor-core tile (6 and 16 output channels, K = 25 and 150), so the
program is the "small" path of SURVEY 2.4: im2col + the tcgen05 GEMM with zero-padded operands.
Channel counts that are not multiples of 8 are padded *physically* (6 -> 8 channels in the activations
and in the packed filter copies; 84 -> 88 columns in the fc operands) because TMA wants 16-byte row
strides; the padding lanes hold zeros and never reach the parameters, whose names, shapes and
values are exactly the reference's `state_dict` (conv weights are stored [Cout][KH][KW][Cin] in the
arena and exposed as torch's [Cout][Cin][KH][KW] views, like every native model).
Operand copies are re-packed from the fp32 masters each step (44 K parameters), so the model trains
correctly under any optimizer, including the stock `torch.optim.SGD` the reference's function uses.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from .. import ops
from ..ops import nn as K
from .native import Buffers, NativeNet, NConv2d, NLinear, act_dtype


def _c8(c: int) -> int:
    return (c + 7) // 8 * 8


class NativeLeNet(NativeNet):
    def __init__(self, num_classes: int = 10):
        super().__init__()
        self.conv1 = NConv2d(1, 6, 5, bias=True)
        self.conv2 = NConv2d(6, 16, 5, bias=True)
        self.fc1 = NLinear(256, 120)
        self.fc2 = NLinear(120, 84)
        self.fc3 = NLinear(84, num_classes)
        # torch's default initialisation for Conv2d / Linear (kaiming_uniform with a = sqrt(5))
        for m in (self.conv1, self.conv2):
            nn.init.kaiming_uniform_(m.weight, a=5 ** 0.5)
            bound = 1.0 / (m.cin * m.k * m.k) ** 0.5
            nn.init.uniform_(m.bias, -bound, bound)
        for m in (self.fc1, self.fc2, self.fc3):
            nn.init.kaiming_uniform_(m.weight, a=5 ** 0.5)
        self.num_classes = num_classes
        self._saved: Optional[dict] = None

    # ------------------------------------------------------------------------------------ helpers
    @staticmethod
    def _pack_filter(conv: NConv2d, B: Buffers, tag: str, cin_phys: int) -> torch.Tensor:
        """bf16 [Cout, ld] filter rows in (kh, kw, c_phys) order, zero in the padded channels/columns."""
        kd = conv.k * conv.k * cin_phys
        ld = (kd + 63) // 64 * 64
        wp = B.get(tag + ".wpack", (conv.cout, ld), zero=True)
        src = conv._w_master if getattr(conv, "_w_master", None) is not None else conv.weight.detach().permute(0, 2, 3, 1)
        wp[:, :kd].view(conv.cout, conv.k * conv.k, cin_phys)[:, :, :conv.cin].copy_(
            src.reshape(conv.cout, conv.k * conv.k, conv.cin))
        return wp

    @staticmethod
    def _pack_fc(fc: NLinear, B: Buffers, tag: str) -> torch.Tensor:
        """bf16 [cout, cin_padded] view-compatible copy of an fc weight whose row stride TMA accepts."""
        ld = _c8(fc.cin)
        wp = B.get(tag + ".wpack", (fc.cout, ld), zero=True)
        src = fc._w_master if getattr(fc, "_w_master", None) is not None else fc.weight.detach()
        wp[:, :fc.cin].copy_(src)
        return wp

    def _conv(self, conv: NConv2d, x: torch.Tensor, w: torch.Tensor, B: Buffers, tag: str):
        """relu(conv(x) + b) on an activation with physically padded channels; returns (y, cols)."""
        N, H, W, Cp = x.shape
        Ho, Wo = H - conv.k + 1, W - conv.k + 1
        cols = B.get(tag + ".cols", (N * Ho * Wo, w.shape[1]), zero=True)
        K.im2col(x, cols, conv.k, conv.k, 1, 0)
        cp = _c8(conv.cout)
        y = B.get(tag + ".y", (N, Ho, Wo, cp), zero=True)            # padded channels stay zero
        ops.gemm_bf16(cols, w, bias=conv.bias.data, relu=True, out=y.view(-1, cp)[:, :conv.cout], block_n=64)
        return y, cols

    @staticmethod
    def _pool(x: torch.Tensor, B: Buffers, tag: str):
        N, H, W, C = x.shape
        y = B.get(tag + ".y", (N, H // 2, W // 2, C))
        idx = B.get(tag + ".idx", (N, H // 2, W // 2, C), torch.uint8)
        K.maxpool_fwd(x, y, idx, 2, 2, 0)
        return y

    def _fc(self, fc: NLinear, x: torch.Tensor, w: torch.Tensor, B: Buffers, tag: str) -> torch.Tensor:
        ld = _c8(fc.cout)
        y = B.get(tag + ".y", (x.shape[0], ld), zero=True)
        ops.gemm_bf16(x, w[:, :fc.cin], bias=fc.bias.data, relu=True, out=y[:, :fc.cout], block_n=64)
        return y                                                      # padded, zeros beyond cout

    # ------------------------------------------------------------------------------------ forward
    def _forward_impl(self, x: torch.Tensor, train: bool) -> torch.Tensor:
        if x.dim() == 4 and x.shape[1] == 1:                          # NCHW with one channel == NHWC
            x = x.reshape(x.shape[0], x.shape[2], x.shape[3], 1)
        elif x.dim() == 3:
            x = x.unsqueeze(-1)
        x = x.contiguous().to(act_dtype())
        N = x.shape[0]
        B = self.buffers_for(N, train)
        w1 = self._pack_filter(self.conv1, B, "c1", 1)
        w2 = self._pack_filter(self.conv2, B, "c2", 8)
        f1, f2, f3 = (self._pack_fc(m, B, t) for m, t in ((self.fc1, "f1"), (self.fc2, "f2"), (self.fc3, "f3")))
        y1, cols1 = self._conv(self.conv1, x, w1, B, "c1")           # [N,24,24,8]
        p1 = self._pool(y1, B, "p1")                                 # [N,12,12,8]
        y2, cols2 = self._conv(self.conv2, p1, w2, B, "c2")          # [N,8,8,16]
        p2 = self._pool(y2, B, "p2")                                 # [N,4,4,16]
        flat = B.get("flat", (N, 256))
        flat.view(N, 16, 4, 4).copy_(p2.permute(0, 3, 1, 2))          # the reference flattens NCHW
        a1 = self._fc(self.fc1, flat, f1, B, "f1")                   # [N,120]
        a2 = self._fc(self.fc2, a1[:, :120], f2, B, "f2")            # [N,88] (84 valid)
        a3 = self._fc(self.fc3, a2[:, :84], f3, B, "f3")             # [N,16] (classes valid)
        if train:
            self._saved = dict(B=B, x=x, cols1=cols1, y1=y1, p1=p1, cols2=cols2, y2=y2, p2=p2, flat=flat,
                               a1=a1, a2=a2, a3=a3, w2=w2, f1=f1, f2=f2, f3=f3)
        return a3[:, :self.num_classes]

    # ------------------------------------------------------------------------------------ backward
    def _fc_backward(self, fc: NLinear, x: torch.Tensor, y: torch.Tensor, dy: torch.Tensor, w: torch.Tensor,
                     B: Buffers, tag: str, need_dx: bool = True):
        """x: [N, cin] view, y: padded output (ReLU mask), dy: padded gradient [N, ld]."""
        g = B.get(tag + ".g", tuple(y.shape), zero=True)
        K.relu_bwd(dy, y, g, 1.0)                                     # padded columns: y == 0 -> g == 0
        gv = g[:, :fc.cout]
        ops.gemm_bf16(gv, x, a_mn=True, b_mn=True, out_dtype=torch.float32, out=fc.w_grad, block_n=64)
        K.colsum(gv, fc.b_grad)
        if not need_dx:
            return None
        ld = _c8(fc.cin)
        dx = B.get(tag + ".dx", (x.shape[0], ld), zero=True)
        ops.gemm_bf16(gv, w[:, :fc.cin], b_mn=True, out=dx[:, :fc.cin], block_n=64)
        return dx

    def _conv_backward(self, conv: NConv2d, cols: torch.Tensor, y: torch.Tensor, dy: torch.Tensor, w: torch.Tensor,
                       B: Buffers, tag: str, cin_phys: int, in_shape=None):
        cp = y.shape[-1]
        g = B.get(tag + ".g", tuple(y.shape), zero=True)
        bsum = B.get(tag + ".bsum", (cp,), torch.float32)
        K.relu_bwd_cols(dy, y, g, 1.0, bsum)
        conv.b_grad.copy_(bsum[:conv.cout])
        kd = conv.k * conv.k * cin_phys
        g2 = g.view(-1, cp)
        dwp = B.get(tag + ".dw", (cp, cols.shape[1]), torch.float32)
        # the reduction runs over every output pixel of the batch: split it over the GPU (red.add into zeros)
        splits = max(1, min(32, (g2.shape[0] // 64) // 8))
        if splits > 1:
            ops.fill_(dwp.view(-1), 0.0)
        ops.gemm_bf16(g2, cols, a_mn=True, b_mn=True, out_dtype=torch.float32, out=dwp, block_n=64, splits=splits)
        conv.w_grad.view(conv.cout, conv.k * conv.k, conv.cin).copy_(
            dwp[:conv.cout, :kd].view(conv.cout, conv.k * conv.k, cin_phys)[:, :, :conv.cin])
        if in_shape is None:
            return None
        dcols = B.get(tag + ".dcols", tuple(cols.shape), zero=True)
        ops.gemm_bf16(g2[:, :conv.cout], w[:, :kd], b_mn=True, out=dcols[:, :kd], block_n=64)
        dx = B.get(tag + ".dx", tuple(in_shape))
        K.col2im(dcols, None, dx, conv.k, conv.k, 1, 0)
        return dx

    def _backward_impl(self, dlogits: torch.Tensor) -> None:
        S = self._saved
        B: Buffers = S["B"]
        N = dlogits.shape[0]
        d3 = B.get("f3.dy", tuple(S["a3"].shape), zero=True)
        d3[:, :self.num_classes].copy_(dlogits)
        d2 = self._fc_backward(self.fc3, S["a2"][:, :84], S["a3"], d3, S["f3"], B, "f3")
        d1 = self._fc_backward(self.fc2, S["a1"][:, :120], S["a2"], d2, S["f2"], B, "f2")
        dflat = self._fc_backward(self.fc1, S["flat"], S["a1"], d1, S["f1"], B, "f1")
        dp2 = B.get("p2.dy", tuple(S["p2"].shape))
        dp2.copy_(dflat[:, :256].view(N, 16, 4, 4).permute(0, 2, 3, 1))
        dy2 = B.get("p2.dx", tuple(S["y2"].shape))
        K.maxpool_bwd(dp2, B["p2.idx"], dy2, 2, 2, 0)
        dp1 = self._conv_backward(self.conv2, S["cols2"], S["y2"], dy2, S["w2"], B, "c2", 8, in_shape=S["p1"].shape)
        dy1 = B.get("p1.dx", tuple(S["y1"].shape))
        K.maxpool_bwd(dp1, B["p1.idx"], dy1, 2, 2, 0)
        self._conv_backward(self.conv1, S["cols1"], S["y1"], dy1, None, B, "c1", 1)
        self._rebind_grads()


def lenet(num_classes: int = 10) -> NativeLeNet:
    return NativeLeNet(num_classes)
