"""Building blocks of the native (hand-written-kernel) models.

A native network is an `nn.Module` whose parameters carry torchvision's names and logical shapes
(so `state_dict()`, checkpoints and the K-AVG merge see exactly what the reference's functions
publish, python/kubeml/kubeml/network.py:444-461) but whose forward and backward are an explicit
program of kernel launches over NHWC bf16 activations:

* parameters are fp32 masters in the `WeightArena`; convolution filters are stored physically as
  `[Cout][KH][KW][Cin]` (the parameter is the permuted view), their bf16 tensor-core copies come from
  the optimizer / merge kernels (never a separate cast pass per layer);
* autograd sees the whole network as ONE function: `forward(x) -> logits`; its backward runs the
  hand-ordered gradient program and writes parameter gradients straight into the arena's flat
  gradient buffer, so no per-op autograd bookkeeping exists on the hot path and the step can be
  captured in a CUDA graph;
* every activation / workspace buffer is allocated once per batch size and reused.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from .. import ops
from ..ops import nn as K


# Activation dtype.  bf16 on the GPU (the kernels accept nothing else); the CPU restatement can be
# run in fp32 (tests) to check the forward/backward *program* against torchvision without rounding.
ACT_DTYPE = torch.bfloat16


def set_activation_dtype(dtype: torch.dtype) -> None:
    global ACT_DTYPE
    ACT_DTYPE = dtype


def act_dtype() -> torch.dtype:
    return ACT_DTYPE


def _ceil(a: int, b: int) -> int:
    return (a + b - 1) // b * b


class NConv2d(nn.Module):
    """Parameter holder with torch's `Conv2d` naming: `weight` [Cout,Cin,KH,KW] (+ `bias`)."""

    def __init__(self, cin: int, cout: int, k: int, stride: int = 1, padding: int = 0, bias: bool = False):
        super().__init__()
        self.cin, self.cout, self.k, self.stride, self.padding = cin, cout, k, stride, padding
        w = torch.empty(cout, cin, k, k)
        nn.init.kaiming_normal_(w, mode="fan_out", nonlinearity="relu")
        self.weight = nn.Parameter(w)
        self.weight._kml_layout = "OHWI"
        self.bias = nn.Parameter(torch.zeros(cout)) if bias else None
        # set by the owning network when the arena is bound
        self.w_bf16: Optional[torch.Tensor] = None      # [Cout,KH,KW,Cin] bf16 (arena slice)
        self.w_grad: Optional[torch.Tensor] = None      # [Cout,KH,KW,Cin] fp32 (arena slice)
        self.b_grad: Optional[torch.Tensor] = None

    @property
    def direct(self) -> bool:
        """Can the shifted-TMA implicit GEMM take this layer (else im2col + GEMM)?"""
        return self.stride == 1 and self.cin % 64 == 0 and self.cout % 8 == 0


class NBatchNorm2d(nn.Module):
    def __init__(self, c: int, eps: float = 1e-5, momentum: float = 0.1):
        super().__init__()
        self.c, self.eps, self.momentum = c, eps, momentum
        self.weight = nn.Parameter(torch.ones(c))
        self.bias = nn.Parameter(torch.zeros(c))
        self.register_buffer("running_mean", torch.zeros(c))
        self.register_buffer("running_var", torch.ones(c))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))
        self.g_weight: Optional[torch.Tensor] = None
        self.g_bias: Optional[torch.Tensor] = None


class NLinear(nn.Module):
    def __init__(self, cin: int, cout: int):
        super().__init__()
        self.cin, self.cout = cin, cout
        bound = 1.0 / cin ** 0.5
        self.weight = nn.Parameter(torch.empty(cout, cin).uniform_(-bound, bound))
        self.bias = nn.Parameter(torch.empty(cout).uniform_(-bound, bound))
        self.w_bf16 = self.w_grad = self.b_grad = None


class Buffers:
    """Named tensors allocated once per (batch size, mode) and reused by every step."""

    def __init__(self, device: torch.device):
        self.device = device
        self._t: Dict[str, torch.Tensor] = {}

    def get(self, name: str, shape: Tuple[int, ...], dtype=None, zero: bool = False) -> torch.Tensor:
        dtype = dtype or ACT_DTYPE
        t = self._t.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = (torch.zeros if zero else torch.empty)(shape, dtype=dtype, device=self.device)
            self._t[name] = t
        return t

    def __getitem__(self, name: str) -> torch.Tensor:
        return self._t[name]

    def has(self, name: str) -> bool:
        return name in self._t


class _NetFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, anchor, net):
        ctx.net = net
        return net._forward_impl(x, train=True)

    @staticmethod
    def backward(ctx, dlogits):
        ctx.net._backward_impl(dlogits.contiguous())
        # The anchor must receive a *defined* gradient: with an undefined one autograd's engine
        # mishandles the leaf's stream bookkeeping under CUDA-graph capture ("dependency created on
        # uncaptured work in another stream").  Parameter gradients themselves are written straight
        # into the arena by the backward program above.
        return None, ctx.net._anchor_zero, None


class NativeNet(nn.Module):
    """Base class: arena hooks, the one-function autograd bridge, buffer cache."""

    native = True

    def __init__(self):
        super().__init__()
        self._arena = None
        self._bufs: Dict[Tuple, Buffers] = {}
        self._grad_zero_range: Optional[Tuple[int, int]] = None

    # -- arena -----------------------------------------------------------------------------
    def _on_arena_bound(self, arena) -> None:
        self._arena = arena
        names = {id(p): n for n, p in self.named_parameters()}
        for m in self.modules():
            if isinstance(m, (NConv2d, NLinear)):
                wn = names[id(m.weight)]
                m.w_bf16 = arena.physical(wn, "bf16")
                m.w_grad = arena.physical(wn, "grads")
                m._w_master = arena.physical(wn, "weights")
                if m.bias is not None:
                    m.b_grad = arena.physical(names[id(m.bias)], "grads")
            elif isinstance(m, NBatchNorm2d):
                m.g_weight = arena.physical(names[id(m.weight)], "grads")
                m.g_bias = arena.physical(names[id(m.bias)], "grads")
        self._after_bound(arena)

    def _after_bound(self, arena) -> None:
        pass

    def _ensure_arena(self) -> None:
        """Stand-alone use (no train job): give the network a private arena on its device."""
        if self._arena is None:
            from ..parallel.arena import ArenaLayout, WeightArena
            dev = next(self.parameters()).device
            arena = WeightArena(ArenaLayout(self), dev, exportable=False)
            arena.bind(self, copy_in=True)

    def buffers_for(self, n: int, train: bool) -> Buffers:
        key = (n, train)
        b = self._bufs.get(key)
        if b is None:
            b = Buffers(next(self.parameters()).device)
            self._bufs[key] = b
        return b

    # -- nn.Module surface -----------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        self._ensure_arena()
        if self.training and torch.is_grad_enabled():
            return _NetFunction.apply(x, self._anchor(), self)
        return self._forward_impl(x, train=self.training)

    def _anchor(self) -> torch.Tensor:
        """A fresh requires-grad scalar per forward that ties the one-function graph to autograd.
        Fresh on purpose: autograd pins a leaf's AccumulateGrad node to the stream that was current
        when the node was created, so a leaf made during eager warm-up would drag the warm-up stream
        into a later CUDA-graph capture."""
        dev = next(self.parameters()).device
        z = getattr(self, "_anchor_zero", None)
        if z is None or z.device != dev:
            self._anchor_zero = torch.zeros(1, device=dev)
        return torch.empty(1, device=dev, requires_grad=True)

    def _forward_impl(self, x: torch.Tensor, train: bool) -> torch.Tensor:
        raise NotImplementedError

    def _backward_impl(self, dlogits: torch.Tensor) -> None:
        raise NotImplementedError

    def _apply(self, fn, recurse=True):
        # moving the module (.to(device)) invalidates arena slices and cached buffers
        out = super()._apply(fn, recurse)
        return out


# ---------------------------------------------------------------------------------------------
# layer programs shared by the models
# ---------------------------------------------------------------------------------------------
def conv_forward(conv: NConv2d, x: torch.Tensor, B: Buffers, tag: str, bias=None, relu: bool = False,
                 splits: int = 1) -> torch.Tensor:
    """x: bf16 [N,H,W,Cin] -> bf16 [N,Ho,Wo,Cout]; saves im2col columns under `tag` when used."""
    N, H, W, _ = x.shape
    k, s, p = conv.k, conv.stride, conv.padding
    Ho, Wo = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
    y = B.get(tag + ".y", (N, Ho, Wo, conv.cout))
    if conv.direct and ops.conv2d_fprop_supported(H, W, conv.cin, conv.cout, k, k, p):
        ops.conv2d_fprop(x, conv.w_bf16, p, bias=bias, relu=relu, out=y)
    else:
        Kd = k * k * conv.cin
        ldk = _ceil(Kd, 64)
        cols = B.get(tag + ".cols", (N * Ho * Wo, ldk))
        K.im2col(x, cols, k, k, s, p)
        ops.gemm_bf16(cols, _packed_weight(conv, B, tag, ldk), bias=bias, relu=relu,
                      out=y.view(N * Ho * Wo, conv.cout))
    return y


def _packed_weight(conv: NConv2d, B: Buffers, tag: str, ldk: int) -> torch.Tensor:
    """[Cout, ldk] bf16 view of the filter; a padded copy only when K is not a multiple of 8."""
    Kd = conv.k * conv.k * conv.cin
    if Kd % 8 == 0:
        return conv.w_bf16.view(conv.cout, Kd)
    wp = B.get(tag + ".wpack", (conv.cout, ldk))
    K.pack_rows(conv._w_master.view(conv.cout, Kd), wp)
    return wp


def conv_backward(conv: NConv2d, x: torch.Tensor, dy: torch.Tensor, B: Buffers, tag: str,
                  need_dx: bool = True, addend: Optional[torch.Tensor] = None, wgrad_splits: int = 1,
                  dx_out: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """Weight gradient into the arena, data gradient (+ addend) returned."""
    N, H, W, _ = x.shape
    k, s, p = conv.k, conv.stride, conv.padding
    Ho, Wo = dy.shape[1], dy.shape[2]
    M = N * Ho * Wo
    dy2 = dy.view(M, conv.cout)
    direct = conv.direct and ops.conv2d_fprop_supported(H, W, conv.cin, conv.cout, k, k, p)
    if direct:
        ops.conv2d_wgrad(x, dy, k, k, p, out=conv.w_grad, splits=wgrad_splits)
        if not need_dx:
            return None
        dx = dx_out if dx_out is not None else B.get(tag + ".dx", (N, H, W, conv.cin))
        ops.conv2d_dgrad(dy, conv.w_bf16, p, H, W, addend=addend, out=dx)
        return dx
    Kd = k * k * conv.cin
    cols = B[tag + ".cols"]
    ldk = cols.shape[1]
    # dW[Cout][Kd] = dY^T · cols  (both consumed MN-major, reduction over the M output pixels)
    ops.gemm_bf16(dy2, cols[:, :Kd] if Kd % 8 else cols, a_mn=True, b_mn=True, out_dtype=torch.float32,
                  out=conv.w_grad.view(conv.cout, Kd), splits=wgrad_splits)
    if not need_dx:
        return None
    # dcols[M][Kd] = dY · W  (W consumed MN-major: no transposed weight copy)
    dcols = B.get(tag + ".dcols", (M, ldk))
    ops.gemm_bf16(dy2, conv.w_bf16.view(conv.cout, Kd), b_mn=True, out=dcols[:, :Kd] if ldk != Kd else dcols)
    dx = dx_out if dx_out is not None else B.get(tag + ".dx", (N, H, W, conv.cin))
    K.col2im(dcols, addend, dx, k, k, s, p)
    return dx


def bn_forward(bn: NBatchNorm2d, x: torch.Tensor, B: Buffers, tag: str, train: bool, relu: bool,
               res: Optional[torch.Tensor] = None) -> torch.Tensor:
    y = B.get(tag + ".y", tuple(x.shape))
    if train:
        mean = B.get(tag + ".mean", (bn.c,), torch.float32)
        istd = B.get(tag + ".istd", (bn.c,), torch.float32)
        K.bn_fwd_train(x, res, y, bn.weight.data, bn.bias.data, bn.running_mean, bn.running_var,
                       bn.num_batches_tracked, mean, istd, relu, bn.eps, bn.momentum)
    else:
        K.bn_fwd_eval(x, res, y, bn.weight.data, bn.bias.data, bn.running_mean, bn.running_var, relu, bn.eps)
    return y


def bn_backward(bn: NBatchNorm2d, dy: torch.Tensor, y: Optional[torch.Tensor], x: torch.Tensor, B: Buffers,
                tag: str, want_dres: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    dx = B.get(tag + ".dx", tuple(x.shape))
    dres = B.get(tag + ".dres", tuple(x.shape)) if want_dres else None
    K.bn_bwd(dy, y, x, bn.weight.data, B[tag + ".mean"], B[tag + ".istd"], dx, dres, bn.g_weight, bn.g_bias)
    return dx, dres


def linear_forward(fc: NLinear, x: torch.Tensor, B: Buffers, tag: str, relu: bool = False) -> torch.Tensor:
    """y[N, cout]; the row stride is padded to a multiple of 8 elements (TMA wants 16-byte strides
    when the gradient GEMMs read it back), the returned tensor is the [:, :cout] view."""
    ld = _ceil(fc.cout, 8)
    y = B.get(tag + ".y", (x.shape[0], ld))
    ops.gemm_bf16(x, fc.w_bf16, bias=fc.bias.data, relu=relu, out=y[:, :fc.cout] if ld != fc.cout else y)
    return y[:, :fc.cout] if ld != fc.cout else y


def linear_backward(fc: NLinear, x: torch.Tensor, dy: torch.Tensor, B: Buffers, tag: str,
                    need_dx: bool = True) -> Optional[torch.Tensor]:
    ld = _ceil(fc.cout, 8)
    if ld != fc.cout or dy.stride(0) != ld or not dy.is_contiguous():
        pad = B.get(tag + ".dy", (dy.shape[0], ld), zero=True)
        pad[:, :fc.cout].copy_(dy)
        dy = pad[:, :fc.cout]
    ops.gemm_bf16(dy, x, a_mn=True, b_mn=True, out_dtype=torch.float32, out=fc.w_grad)     # dW = dY^T · X
    K.colsum(dy, fc.b_grad)
    if not need_dx:
        return None
    dx = B.get(tag + ".dx", (x.shape[0], fc.cin))
    ops.gemm_bf16(dy, fc.w_bf16, b_mn=True, out=dx)                                          # dX = dY · W
    return dx
