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

import os
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


def live_taps(H: int, W: int, Ho: int, Wo: int, k: int, p: int, s: int = 1) -> int:
    """Filter taps that read at least one real pixel (the others lie entirely in the zero padding and are
    skipped by the implicit GEMM): 1 of 9 on a 1x1 map, 4 of 9 for a 3x3/2 convolution from 2x2 to 1x1."""
    def hit(kk, extent, outs):
        return any(0 <= s * o + kk - p < extent for o in range(outs))
    return sum(1 for kh in range(k) for kw in range(k) if hit(kh, H, Ho) and hit(kw, W, Wo))


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
        """Can the shifted-TMA implicit GEMM take this layer (else im2col + GEMM)?  Any stride: the tensor
        map walks the input with an element stride.  (Only the data gradient of a strided layer still goes
        through a column buffer.)"""
        return self.cin % 64 == 0 and self.cout % 8 == 0


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
        self.split_grads: List[torch.Tensor] = []     # gradients accumulated with red.add this pass
        self.zero_range: Optional[Tuple[int, int]] = None

    def get(self, name: str, shape: Tuple[int, ...], dtype=None, zero: bool = False) -> torch.Tensor:
        dtype = dtype or ACT_DTYPE
        t = self._t.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = (torch.zeros if zero else torch.empty)(shape, dtype=dtype, device=self.device)
            self._t[name] = t
        return t

    def __getitem__(self, name: str) -> torch.Tensor:
        return self._t[name]

    # -- side stream: work that is off the critical path (weight gradients) runs next to the
    #    data-gradient chain; both are latency-bound kernels on a handful of SMs each -------------
    def on_side(self):
        import contextlib
        # measured on B200: a loss while BatchNorm used cooperative launches (which cannot overlap
        # anything: 3.04 vs 2.72 ms per ResNet-34 step), a gain since it uses a software grid barrier
        # and launches are programmatic-dependent (1.82 vs 1.95 ms)
        if self.device.type != "cuda" or os.environ.get("KUBEML_SIDE_STREAM", "1") == "0":
            return contextlib.nullcontext()
        if getattr(self, "_side", None) is None:
            self._side = torch.cuda.Stream(device=self.device)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._side.wait_event(ev)
        self._side_used = True
        return torch.cuda.stream(self._side)

    def join_side(self) -> None:
        if getattr(self, "_side_used", False):
            torch.cuda.current_stream(self.device).wait_stream(self._side)
            self._side_used = False

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

    def _begin_backward(self, B: "Buffers") -> None:
        """Split-K weight gradients are accumulated with red.add and must start from zero.  The first
        backward of a batch size clears the whole gradient buffer and records which tensors were
        accumulated; later ones clear only the flat range that covers those."""
        a = self._arena
        n = a.layout.n_param
        if B.zero_range is None:
            ops.fill_(a.grads[:n], 0.0)
        elif B.zero_range[1] > B.zero_range[0]:
            ops.fill_(a.grads[B.zero_range[0]:B.zero_range[1]], 0.0)
        B.split_grads = []

    def _rebind_grads(self) -> None:
        """A stock optimizer's `zero_grad()` drops `.grad` (set_to_none): point it back at the arena
        slices the backward program wrote."""
        first = next(self.parameters())
        if first.grad is not None and first.grad.data_ptr() == self._arena.grads.data_ptr():
            return
        for name, p in self.named_parameters():
            p.grad = self._arena.grad(name)

    def _end_backward(self, B: "Buffers") -> None:
        B.join_side()
        self._rebind_grads()
        if B.zero_range is None:
            base = self._arena.grads.data_ptr()
            lo, hi = 0, 0
            if B.split_grads:
                lo = min((t.data_ptr() - base) // 4 for t in B.split_grads)
                hi = max((t.data_ptr() - base) // 4 + t.numel() for t in B.split_grads)
                lo, hi = lo // 4 * 4, (hi + 3) // 4 * 4
            B.zero_range = (lo, hi)

    def _forward_impl(self, x: torch.Tensor, train: bool) -> torch.Tensor:
        raise NotImplementedError

    def _backward_impl(self, dlogits: torch.Tensor) -> None:
        raise NotImplementedError

    def load_state_dict(self, *args, **kwargs):
        """Host-path syncs (TensorStore / collectives) load merged weights through here: the
        bf16 tensor-core copy has to follow the fp32 masters."""
        out = super().load_state_dict(*args, **kwargs)
        if self._arena is not None:
            self._arena.refresh_bf16()
        return out

    def _apply(self, fn, recurse=True):
        # moving the module (.to(device)) invalidates arena slices and cached buffers
        out = super()._apply(fn, recurse)
        return out


# ---------------------------------------------------------------------------------------------
# layer programs shared by the models
# ---------------------------------------------------------------------------------------------
_sms_cached: Optional[int] = None


def sm_count() -> int:
    """SMs of the GPU the plans are sized for: asked from the native library (`kml_num_sms`) when a device is
    present, 148 (B200) for the CPU restatement."""
    global _sms_cached
    if _sms_cached is None:
        n = 148
        if torch.cuda.is_available():
            try:
                from .. import _native
                n = int(_native.lib().kml_num_sms()) or 148
            except Exception:  # noqa: BLE001 - planning must not fail because the count is unknown
                n = 148
        _sms_cached = n
    return _sms_cached


def auto_splits(tiles: int, chunks: int, min_chunks: int = 2, budget: Optional[int] = None) -> int:
    """K-splits for a GEMM with `tiles` output tiles and `chunks` 64-deep reduction chunks: enough
    CTAs to cover the GPU (`budget` CTAs), never fewer than `min_chunks` chunks per CTA, and only
    split counts that leave no CTA without work."""
    if budget is None:
        budget = sm_count()
    if tiles >= budget or chunks < 2 * min_chunks:
        return 1
    s = max(1, min(budget // max(tiles, 1), chunks // min_chunks))
    per = (chunks + s - 1) // s
    return (chunks + per - 1) // per


def pick_block_n(m_tiles: int, n: int) -> int:
    """128-wide tiles only when there are enough of them to occupy the GPU."""
    if n <= 64:
        return 64
    return 128 if m_tiles * ((n + 127) // 128) >= sm_count() // 2 else 64


def plan_cluster(m_tiles: int, n: int, chunks: int, bn0: int, base_budget: Optional[int] = None):
    """(block_n, splits) for a cluster split-K GEMM.  In this regime a CTA's time is its number of K
    chunks times a TMA round trip, whatever the tile width (measured: 0.42 us per chunk for 128x64 and
    128x128 alike), so when the default choice leaves long K loops a wider tile - half the tiles, twice
    the splits - halves the time (VGG 512->512 on 4x4: 72 -> 36 chunks per CTA)."""
    def plan(bn, budget=None):
        budget = budget or base_budget or sm_count()
        tiles = m_tiles * ((n + bn - 1) // bn)
        S = min(auto_splits(tiles, chunks, min_chunks=_CL_MIN_CHUNKS, budget=budget), _CLUSTER_MAX)
        return -(-chunks // S), S, bn
    best = plan(bn0)
    if best[0] > 12:                       # long K loops: consider wider tiles and two CTAs per SM
        cands = [plan(bn0, 2 * sm_count())]
        if bn0 == 64 and n >= 128:
            cands += [plan(128), plan(128, 2 * sm_count())]
        for c in cands:
            if c[0] < best[0]:
                best = c
    return best[2], best[1]


class Act:
    """An activation or gradient: a bf16 tensor, or the fp32 split-K *slabs* [S, ...] of the GEMM that
    produced it.  Slabs are folded by the BatchNorm kernel that consumes them (csrc/norm.cu), so a
    small, deep GEMM can be spread over the whole GPU without a reduction pass or atomics.
    `stats`: per-channel sum / sum of squares the producing GEMM accumulated (cluster split-K)."""
    __slots__ = ("t", "slabs", "stats", "bn_done", "g")

    def __init__(self, t: torch.Tensor, slabs: int = 0, stats: Optional[torch.Tensor] = None,
                 bn_done: bool = False, g: Optional[torch.Tensor] = None):
        self.t, self.slabs, self.stats = t, slabs, stats
        # bn_done: the producing data-gradient GEMM already ran the consuming BatchNorm's backward
        # (kml_bn_fuse mode 2): `t` is that layer's dx, `g` its masked incoming gradient (if asked for)
        self.bn_done, self.g = bn_done, g

    @property
    def shape(self):
        return self.t.shape[1:] if self.slabs else self.t.shape

    def dense(self, B: "Buffers", tag: str) -> torch.Tensor:
        """bf16 tensor (folds slabs with a torch op; only used off the hot path)."""
        if not self.slabs:
            return self.t
        out = B.get(tag + ".dense", tuple(self.shape))
        out.copy_(self.t.sum(0))
        return out


def _slab_enabled() -> bool:
    return os.environ.get("KUBEML_SPLITK_SLABS", "1") != "0"


_CLUSTER_MAX = int(os.environ.get("KUBEML_CLUSTER_MAX", "8"))       # 8 = portable cluster size, 16 = B200 maximum
_CL_MIN_CHUNKS = int(os.environ.get("KUBEML_CL_MIN_CHUNKS", "1"))   # fewest K chunks per CTA of a cluster split


def _bn_bwd_fusion() -> bool:
    """Let the data-gradient GEMM also mask and pre-reduce for the BatchNorm backward behind it.
    Off by default: measured neutral on B200 (ResNet-34 step 1.585 ms without, 1.594 ms with) - the
    heavier GEMM epilogue costs what the barrier-free BatchNorm kernel saves."""
    return os.environ.get("KUBEML_CLUSTER_BNBWD", "0") == "1"


def _bn_fuse_enabled() -> bool:
    """Finish BatchNorm inside the GEMM that feeds it (forward: conv + BN + residual + ReLU in one launch;
    backward: data gradient + BN backward in one launch) - csrc/gemm_tcgen05.cu `kml_bn_fuse`.  The GEMM
    then synchronises its whole grid in software, so it is used on the main stream only and only when the
    grid is co-resident (`fused_plan`)."""
    return os.environ.get("KUBEML_FUSE_BN", "1") != "0" and _cluster_enabled()


_coresident_cache: Dict[Tuple, int] = {}


def fused_plan(device: torch.device, m_tiles: int, n: int, chunks: int, bn0: int, b_mn: bool):
    """(block_n, splits) for a GEMM with a fused BatchNorm epilogue, or None when no plan keeps the whole grid
    resident at once (the fused epilogue spins on a grid-wide barrier)."""
    if n > 2048 or n % 8:
        return None
    budget = int(os.environ.get("KUBEML_FUSED_BUDGET", "0")) or None
    bn_, S = plan_cluster(m_tiles, n, chunks, bn0, base_budget=budget)
    while True:
        ctas = m_tiles * ((n + bn_ - 1) // bn_) * S
        if device.type != "cuda":
            return bn_, S
        key = (device.index, bn_, b_mn, S)
        cap = _coresident_cache.get(key)
        if cap is None:
            cap = ops.max_coresident(bn_, False, b_mn, S)
            _coresident_cache[key] = cap
        if 0 < ctas <= cap:
            return bn_, S
        if S == 1:
            return None
        S = max(1, S // 2)
        per = -(-chunks // S)
        S = -(-chunks // per)


def _cluster_enabled() -> bool:
    """Split-K as one thread-block cluster per tile, reduced through distributed shared memory in the
    GEMM itself (final bf16 output + BatchNorm statistics) instead of fp32 slabs folded by the consumer."""
    return os.environ.get("KUBEML_CLUSTER_SPLITK", "1") != "0"


def conv_forward(conv: NConv2d, x: torch.Tensor, B: Buffers, tag: str, bias=None, relu: bool = False,
                 allow_slabs: bool = False, fold: bool = False, want_stats: bool = False) -> Act:
    """x: bf16 [N,H,W,Cin] -> [N,Ho,Wo,Cout] (bf16, or fp32 slabs when `allow_slabs` and the problem is
    too small to fill the GPU otherwise); saves im2col columns under `tag` when used.  With `fold`
    a split result is reduced here by the bias+activation kernel (no BatchNorm follows to do it)."""
    N, H, W, _ = x.shape
    k, s, p = conv.k, conv.stride, conv.padding
    Ho, Wo = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
    M = N * Ho * Wo
    m_tiles = (M + 127) // 128
    bn_ = pick_block_n(m_tiles, conv.cout)
    tiles = m_tiles * ((conv.cout + bn_ - 1) // bn_)
    direct = conv.direct and ops.conv2d_fprop_supported(H, W, conv.cin, conv.cout, k, k, p, s)
    Kd = k * k * conv.cin
    if direct:
        chunks = live_taps(H, W, Ho, Wo, k, p, s) * (conv.cin // 64)
    else:
        chunks = (Kd + 63) // 64
    if allow_slabs and _cluster_enabled() and (want_stats or fold):
        # cluster split-K: complete output (+ bias/ReLU, + BatchNorm statistics) straight from the GEMM
        bn_, S = plan_cluster(m_tiles, conv.cout, chunks, bn_)
        y = B.get(tag + ".y", (N, Ho, Wo, conv.cout))
        stats = B.get(tag + ".stats", (K.bn_stats_numel(conv.cout),), torch.float32, zero=True) if want_stats else None
        # one split and no statistics wanted: the plain epilogue does the same, and a launch without the
        # cluster attribute may run persistent when it has many tiles
        cl = S > 1 or stats is not None
        if direct:
            ops.conv2d_fprop(x, conv.w_bf16, p, bias=bias, relu=relu, out=y, block_n=bn_, splits=S, cluster=cl,
                             stats=stats, stride=s)
        else:
            ldk = _ceil(Kd, 64)
            cols = B.get(tag + ".cols", (M, ldk))
            K.im2col(x, cols, k, k, s, p)
            ops.gemm_bf16(cols[:, :Kd] if ldk != Kd else cols, _packed_weight(conv, B, tag, ldk)[:, :Kd], bias=bias,
                          relu=relu, out=y.view(M, conv.cout),
                          block_n=bn_, splits=S, cluster=cl, stats=stats)
        return Act(y, 0, stats)
    can_split = allow_slabs and _slab_enabled() and (fold or (bias is None and not relu))
    S = auto_splits(tiles, chunks, min_chunks=3) if can_split else 1
    if S > 1:
        y = B.get(tag + ".slabs", (S, N, Ho, Wo, conv.cout), torch.float32)
        ebias, erelu = None, False
    else:
        y = B.get(tag + ".y", (N, Ho, Wo, conv.cout))
        ebias, erelu = bias, relu
    if direct:
        ops.conv2d_fprop(x, conv.w_bf16, p, bias=ebias, relu=erelu, out=y, block_n=bn_, splits=S, slabs=S > 1, stride=s)
    else:
        ldk = _ceil(Kd, 64)
        cols = B.get(tag + ".cols", (M, ldk))
        K.im2col(x, cols, k, k, s, p)
        ops.gemm_bf16(cols[:, :Kd] if ldk != Kd else cols, _packed_weight(conv, B, tag, ldk)[:, :Kd], bias=ebias, relu=erelu,
                      out=y.view(S, M, conv.cout) if S > 1 else y.view(M, conv.cout), block_n=bn_,
                      splits=S, slabs=S > 1)
    if S > 1 and fold:
        yd = B.get(tag + ".y", (N, Ho, Wo, conv.cout))
        K.bias_act(y.view(S, M, conv.cout), bias, yd.view(M, conv.cout), relu)
        return Act(yd)
    return Act(y, S if S > 1 else 0)


def conv_bn_forward(conv: NConv2d, bn: "NBatchNorm2d", x: torch.Tensor, B: Buffers, tag: str, bntag: str,
                    train: bool, relu: bool, res: Optional[torch.Tensor] = None, fuse: bool = True):
    """relu?(bn(conv(x)) + res?) -> (y, x_saved) like `bn_forward`.  In training, when the GEMM's grid fits the
    GPU at once, convolution and BatchNorm are ONE launch (the GEMM accumulates the statistics, synchronises
    its grid and normalises the tile it still holds); otherwise GEMM (+ statistics) then `bn_forward`."""
    if train and fuse and _bn_fuse_enabled() and conv.bias is None:
        N, H, W, _ = x.shape
        k, s, p = conv.k, conv.stride, conv.padding
        Ho, Wo = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
        M = N * Ho * Wo
        m_tiles = (M + 127) // 128
        direct = conv.direct and ops.conv2d_fprop_supported(H, W, conv.cin, conv.cout, k, k, p, s)
        Kd = k * k * conv.cin
        if direct:
            chunks = live_taps(H, W, Ho, Wo, k, p, s) * (conv.cin // 64)
        else:
            chunks = (Kd + 63) // 64
        plan = fused_plan(x.device, m_tiles, conv.cout, chunks, pick_block_n(m_tiles, conv.cout), False)
        if plan is not None:
            bn_, S = plan
            raw = B.get(tag + ".y", (N, Ho, Wo, conv.cout))
            y = B.get(bntag + ".y", (N, Ho, Wo, conv.cout))
            mean = B.get(bntag + ".mean", (bn.c,), torch.float32)
            istd = B.get(bntag + ".istd", (bn.c,), torch.float32)
            bf = ops.BnFwd(bn.weight.data, bn.bias.data, bn.running_mean, bn.running_var, bn.num_batches_tracked,
                           mean, istd, raw, res=res, relu=relu, eps=bn.eps, momentum=bn.momentum)
            if direct:
                ops.conv2d_fprop(x, conv.w_bf16, p, out=y, block_n=bn_, splits=S, cluster=True, bn_fuse=bf, stride=s)
            else:
                ldk = _ceil(Kd, 64)
                cols = B.get(tag + ".cols", (M, ldk))
                K.im2col(x, cols, k, k, s, p)
                ops.gemm_bf16(cols[:, :Kd] if ldk != Kd else cols, _packed_weight(conv, B, tag, ldk)[:, :Kd],
                              out=y.view(M, conv.cout), block_n=bn_, splits=S, cluster=True, bn_fuse=bf)
            return y, raw
    t = conv_forward(conv, x, B, tag, allow_slabs=True, want_stats=train)
    return bn_forward(bn, t, B, bntag, train, relu=relu, res=res)


def _packed_weight(conv: NConv2d, B: Buffers, tag: str, ldk: int) -> torch.Tensor:
    """[Cout, ldk] bf16 view of the filter; a padded copy only when K is not a multiple of 8."""
    Kd = conv.k * conv.k * conv.cin
    if Kd % 8 == 0:
        return conv.w_bf16.view(conv.cout, Kd)
    wp = B.get(tag + ".wpack", (conv.cout, ldk))
    K.pack_rows(conv._w_master.view(conv.cout, Kd), wp)
    return wp


def conv_backward(conv: NConv2d, x: torch.Tensor, dy: torch.Tensor, B: Buffers, tag: str,
                  need_dx: bool = True, addend: Optional[torch.Tensor] = None,
                  allow_slabs: bool = False, bn_next=None, wgrad_budget: Optional[int] = None) -> Optional[Act]:
    """Weight gradient into the arena; data gradient (+ addend) returned (bf16, or fp32 slabs when
    `allow_slabs`: the consumer must then be a BatchNorm backward).
    `bn_next` = (bn module, tag, y, x_saved) of the BatchNorm whose backward consumes this gradient:
    in cluster mode the GEMM then applies that layer's ReLU mask and accumulates its two reduction
    sums, and the returned Act carries them (`stats`), so the BatchNorm backward is element-wise."""
    N, H, W, _ = x.shape
    k, s, p = conv.k, conv.stride, conv.padding
    Ho, Wo = dy.shape[1], dy.shape[2]
    M = N * Ho * Wo
    dy2 = dy.view(M, conv.cout)
    direct = conv.direct and ops.conv2d_fprop_supported(H, W, conv.cin, conv.cout, k, k, p, s)
    chunks = (M + 63) // 64
    if direct:
        bn_w = 128 if conv.cin > 64 else 64
        tiles = ((conv.cout + 127) // 128) * ((conv.cin + bn_w - 1) // bn_w) * k * k
        # gradients accumulate with red.add across splits; _begin_backward cleared them
        # weight gradients run next to the data-gradient chain on the side stream; where they have slack (the
        # latency-bound ResNet steps) the model aims them at half of the SMs, which leaves the chain more room
        # (measured: 148 -> 74 CTAs, 1.219 -> 1.209 ms per ResNet-34 step; 37 is slower again; VGG, whose
        # weight gradients are large GEMMs, keeps the whole GPU).  KUBEML_WGRAD_BUDGET overrides.
        sp = auto_splits(tiles, chunks, budget=int(os.environ.get("KUBEML_WGRAD_BUDGET", "0")) or wgrad_budget)
        if sp > 1:
            B.split_grads.append(conv.w_grad)
        with B.on_side():
            ops.conv2d_wgrad(x, dy, k, k, p, out=conv.w_grad, splits=sp, block_n=bn_w, stride=s)
        if not need_dx:
            return None
        if s != 1:
            # A strided layer's data gradient scatters: dY . W into a column buffer, then col2im gathers it
            # (+ addend) into dX.  Forward and weight gradient above needed no such buffer.
            Kd = k * k * conv.cin
            dcols = B.get(tag + ".dcols", (M, Kd))
            ops.gemm_bf16(dy2, conv.w_bf16.view(conv.cout, Kd), b_mn=True, out=dcols,
                          block_n=pick_block_n((M + 127) // 128, Kd))
            dx = B.get(tag + ".dx", (N, H, W, conv.cin))
            K.col2im(dcols, addend, dx, k, k, s, p)
            return Act(dx)
        m_tiles = (N * H * W + 127) // 128
        bn_d = pick_block_n(m_tiles, conv.cin)
        dtiles = m_tiles * ((conv.cin + bn_d - 1) // bn_d)
        pd = k - 1 - p
        live = live_taps(Ho, Wo, H, W, k, pd)
        if allow_slabs and _cluster_enabled():
            bn_d, S = plan_cluster(m_tiles, conv.cin, live * (conv.cout // 64), bn_d)
            if bn_next is not None and _bn_fuse_enabled():
                # data gradient + the consuming BatchNorm's whole backward in one launch
                plan = fused_plan(dy.device, m_tiles, conv.cin, live * (conv.cout // 64), bn_d, True)
                if plan is not None:
                    nbn, ntag, ny, nx = bn_next[:4]
                    want_dres = len(bn_next) > 4 and bool(bn_next[4])
                    bn_d, S = plan
                    out = B.get(ntag + ".dx", (N, H, W, conv.cin))
                    gbuf = B.get(ntag + ".dres", (N, H, W, conv.cin)) if want_dres else None
                    bf = ops.BnBwd(nbn.weight.data, B[ntag + ".mean"], B[ntag + ".istd"], nx, nbn.g_weight, nbn.g_bias,
                                   mask_y=ny, raw_out=gbuf)
                    ops.conv2d_dgrad(dy, conv.w_bf16, p, H, W, addend=addend, out=out, block_n=bn_d, splits=S,
                                     cluster=True, bn_fuse=bf)
                    return Act(out, bn_done=True, g=gbuf)
            dx = B.get(tag + ".dx", (N, H, W, conv.cin))
            if bn_next is not None and _bn_bwd_fusion():
                nbn, ntag, ny, nx = bn_next[:4]
                stats = B.get(ntag + ".bstats", (K.bn_stats_numel(conv.cin),), torch.float32, zero=True)
                ops.conv2d_dgrad(dy, conv.w_bf16, p, H, W, addend=addend, out=dx, block_n=bn_d, splits=S,
                                 cluster=True, mask_y=ny, bn=(nx, B[ntag + ".mean"], B[ntag + ".istd"]), stats=stats)
                return Act(dx, 0, stats)
            ops.conv2d_dgrad(dy, conv.w_bf16, p, H, W, addend=addend, out=dx, block_n=bn_d, splits=S, cluster=S > 1)
            return Act(dx)
        S = auto_splits(dtiles, live * (conv.cout // 64), min_chunks=3) if (allow_slabs and _slab_enabled()) else 1
        if S > 1:
            dx = B.get(tag + ".dxslabs", (S, N, H, W, conv.cin), torch.float32)
        else:
            dx = B.get(tag + ".dx", (N, H, W, conv.cin))
        ops.conv2d_dgrad(dy, conv.w_bf16, p, H, W, addend=addend, out=dx, block_n=bn_d, splits=S, slabs=S > 1)
        return Act(dx, S if S > 1 else 0)
    Kd = k * k * conv.cin
    cols = B[tag + ".cols"]
    ldk = cols.shape[1]
    # dW[Cout][Kd] = dY^T · cols  (both consumed MN-major, reduction over the M output pixels)
    bn_w = 128 if Kd > 64 else 64
    tiles = ((conv.cout + 127) // 128) * ((Kd + bn_w - 1) // bn_w)
    sp = auto_splits(tiles, chunks)
    if sp > 1:
        B.split_grads.append(conv.w_grad)
    with B.on_side():
        ops.gemm_bf16(dy2, cols[:, :Kd] if ldk != Kd else cols, a_mn=True, b_mn=True, out_dtype=torch.float32,
                      out=conv.w_grad.view(conv.cout, Kd), splits=sp, block_n=bn_w)
    if not need_dx:
        return None
    # dcols[M][Kd] = dY · W  (W consumed MN-major: no transposed weight copy)
    dcols = B.get(tag + ".dcols", (M, ldk))
    ops.gemm_bf16(dy2, conv.w_bf16.view(conv.cout, Kd), b_mn=True, out=dcols[:, :Kd] if ldk != Kd else dcols,
                  block_n=pick_block_n((M + 127) // 128, Kd))
    dx = B.get(tag + ".dx", (N, H, W, conv.cin))
    K.col2im(dcols, addend, dx, k, k, s, p)
    return Act(dx)


def bn_forward(bn: NBatchNorm2d, x: Act, B: Buffers, tag: str, train: bool, relu: bool,
               res: Optional[torch.Tensor] = None):
    """Returns (y, x_saved): the block output and the bf16 input the backward pass needs."""
    shape = tuple(x.shape)
    y = B.get(tag + ".y", shape)
    if train and x.stats is not None:
        mean = B.get(tag + ".mean", (bn.c,), torch.float32)
        istd = B.get(tag + ".istd", (bn.c,), torch.float32)
        K.bn_apply_train(x.t, x.stats, res, y, bn.weight.data, bn.bias.data, bn.running_mean, bn.running_var,
                         bn.num_batches_tracked, mean, istd, relu, bn.eps, bn.momentum)
        return y, x.t
    xs = x.t if x.slabs else None
    xb = None if x.slabs else x.t
    x_out = B.get(tag + ".xsave", shape) if x.slabs else None
    if train:
        mean = B.get(tag + ".mean", (bn.c,), torch.float32)
        istd = B.get(tag + ".istd", (bn.c,), torch.float32)
        K.bn_fwd_train(xb, res, y, bn.weight.data, bn.bias.data, bn.running_mean, bn.running_var,
                       bn.num_batches_tracked, mean, istd, relu, bn.eps, bn.momentum, x_slabs=xs, x_out=x_out)
    else:
        K.bn_fwd_eval(xb, res, y, bn.weight.data, bn.bias.data, bn.running_mean, bn.running_var, relu, bn.eps,
                      x_slabs=xs)
    return y, (x_out if x.slabs else x.t)


def bn_backward(bn: NBatchNorm2d, dy: Act, y: Optional[torch.Tensor], x: torch.Tensor, B: Buffers,
                tag: str, want_dres: bool = False):
    if dy.bn_done:                # the data-gradient GEMM already finished this layer's backward
        return dy.t, dy.g
    dx = B.get(tag + ".dx", tuple(x.shape))
    if dy.stats is not None:      # masked gradient + reduction sums came with the data-gradient GEMM
        K.bn_bwd_apply(dy.t, x, bn.weight.data, B[tag + ".mean"], B[tag + ".istd"], dy.stats, dx, bn.g_weight, bn.g_bias)
        return dx, dy.t
    dres = B.get(tag + ".dres", tuple(x.shape)) if want_dres else None
    K.bn_bwd(None if dy.slabs else dy.t, y, x, bn.weight.data, B[tag + ".mean"], B[tag + ".istd"], dx, dres,
             bn.g_weight, bn.g_bias, dy_slabs=dy.t if dy.slabs else None)
    return dx, dres


def linear_forward(fc: NLinear, x: torch.Tensor, B: Buffers, tag: str, relu: bool = False) -> torch.Tensor:
    """y[N, cout]; the row stride is padded to a multiple of 8 elements (TMA wants 16-byte strides
    when the gradient GEMMs read it back), the returned tensor is the [:, :cout] view.  A deep,
    narrow product (VGG's 25088 -> 4096 on 128 rows) is split over K into fp32 slabs that the
    bias+activation kernel folds."""
    ld = _ceil(fc.cout, 8)
    M = x.shape[0]
    y = B.get(tag + ".y", (M, ld))
    yv = y[:, :fc.cout] if ld != fc.cout else y
    m_tiles = (M + 127) // 128
    bn_ = pick_block_n(m_tiles, fc.cout)
    tiles = m_tiles * ((fc.cout + bn_ - 1) // bn_)
    # a long reduction (VGG fc1: 392 chunks) may take two CTAs per SM
    S = auto_splits(tiles, (fc.cin + 63) // 64, min_chunks=8, budget=2 * sm_count()) if (_slab_enabled() and fc.cout % 8 == 0) else 1
    if S > 1:
        slabs = B.get(tag + ".slabs", (S, M, fc.cout), torch.float32)
        ops.gemm_bf16(x, fc.w_bf16, out=slabs, block_n=bn_, splits=S, slabs=True)
        K.bias_act(slabs, fc.bias.data, yv, relu)
    else:
        ops.gemm_bf16(x, fc.w_bf16, bias=fc.bias.data, relu=relu, out=yv, block_n=bn_)
    return yv


def linear_backward(fc: NLinear, x: torch.Tensor, dy: torch.Tensor, B: Buffers, tag: str,
                    need_dx: bool = True, bias_done: bool = False, allow_slabs: bool = False):
    """Weight/bias gradients into the arena (`bias_done`: the caller's mask kernel already produced
    the bias gradient); returns dX as a bf16 tensor, or an `Act` of fp32 split-K slabs when
    `allow_slabs` (the consumer folds them)."""
    ld = _ceil(fc.cout, 8)
    if ld != fc.cout or dy.stride(0) != ld or not dy.is_contiguous():
        pad = B.get(tag + ".dy", (dy.shape[0], ld), zero=True)
        pad[:, :fc.cout].copy_(dy)
        dy = pad[:, :fc.cout]
    with B.on_side():
        ops.gemm_bf16(dy, x, a_mn=True, b_mn=True, out_dtype=torch.float32, out=fc.w_grad)     # dW = dY^T · X (store)
        if not bias_done:
            K.colsum(dy, fc.b_grad)
    if not need_dx:
        return None
    M = x.shape[0]
    m_tiles = (M + 127) // 128
    bn_ = pick_block_n(m_tiles, fc.cin)
    tiles = m_tiles * ((fc.cin + bn_ - 1) // bn_)
    S = auto_splits(tiles, (fc.cout + 63) // 64, min_chunks=8) if (allow_slabs and _slab_enabled()) else 1
    if S > 1:
        dxs = B.get(tag + ".dxslabs", (S, M, fc.cin), torch.float32)
        ops.gemm_bf16(dy, fc.w_bf16, b_mn=True, out=dxs, block_n=bn_, splits=S, slabs=True)   # dX = dY · W
        return Act(dxs, S)
    dx = B.get(tag + ".dx", (M, fc.cin))
    ops.gemm_bf16(dy, fc.w_bf16, b_mn=True, out=dx, block_n=bn_)
    return Act(dx) if allow_slabs else dx
