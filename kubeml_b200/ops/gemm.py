"""Python entry points of the tcgen05 GEMM family (csrc/gemm_tcgen05.cu): plain bf16 GEMMs with
either operand K-major or MN-major, and implicit-GEMM convolution forward / weight gradient on
NHWC bf16 activations.  CPU tensors take an fp32 PyTorch restatement (the test oracle)."""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from .. import _native


def _ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


class BnFwd:
    """A training-mode BatchNorm (+ residual + ReLU) to be finished inside the producing cluster GEMM
    (csrc/gemm_tcgen05.cu, `kml_bn_fuse` mode 1).  `raw_out` receives the GEMM's own output (the
    BatchNorm input, which backward needs); the GEMM's `out` becomes the layer output."""

    def __init__(self, gamma, beta, running_mean, running_var, nbt, save_mean, save_invstd, raw_out,
                 res=None, relu=False, eps=1e-5, momentum=0.1):
        self.gamma, self.beta, self.running_mean, self.running_var, self.nbt = gamma, beta, running_mean, running_var, nbt
        self.save_mean, self.save_invstd, self.raw_out, self.res = save_mean, save_invstd, raw_out, res
        self.relu, self.eps, self.momentum = relu, eps, momentum

    def fill(self, f, device):
        from . import nn as K
        f.mode = 1
        f.scratch = K.bn_scratch(device).data_ptr()
        f.gamma, f.beta = self.gamma.data_ptr(), self.beta.data_ptr()
        f.running_mean, f.running_var = _ptr(self.running_mean) or None, _ptr(self.running_var) or None
        f.num_batches_tracked = _ptr(self.nbt) or None
        f.save_mean, f.save_invstd = self.save_mean.data_ptr(), self.save_invstd.data_ptr()
        f.res, f.raw_out = _ptr(self.res) or None, self.raw_out.data_ptr()
        f.relu, f.eps, f.momentum = int(self.relu), float(self.eps), float(self.momentum)

    def cpu(self, y: torch.Tensor, out: torch.Tensor) -> None:
        """fp32 restatement: y = the GEMM result (already rounded to the output dtype)."""
        C = y.shape[-1]
        M = y.numel() // C
        self.raw_out.copy_(y.reshape(self.raw_out.shape))
        xf = self.raw_out.float().reshape(M, C)
        mean = xf.mean(0)
        var = xf.var(0, unbiased=False)
        invstd = torch.rsqrt(var + self.eps)
        o = (xf - mean) * invstd * self.gamma + self.beta
        if self.res is not None:
            o = o + self.res.float().reshape(M, C)
        if self.relu:
            o = torch.relu(o)
        out.copy_(o.reshape(out.shape))
        self.save_mean.copy_(mean)
        self.save_invstd.copy_(invstd)
        if self.running_mean is not None:
            self.running_mean.mul_(1 - self.momentum).add_(self.momentum * mean)
            self.running_var.mul_(1 - self.momentum).add_(self.momentum * var * (M / max(M - 1, 1)))
        if self.nbt is not None:
            self.nbt.add_(1)


class BnBwd:
    """The backward of a BatchNorm finished inside the data-gradient GEMM that produces its incoming
    gradient (`kml_bn_fuse` mode 2): g = (GEMM result + addend) * (mask_y > 0); the GEMM's `out` becomes
    dx = gamma*invstd*(g - mean(g) - xhat*mean(g*xhat)); `raw_out` (optional) receives g."""

    def __init__(self, gamma, save_mean, save_invstd, bn_x, dgamma, dbeta, mask_y=None, raw_out=None):
        self.gamma, self.save_mean, self.save_invstd, self.bn_x = gamma, save_mean, save_invstd, bn_x
        self.dgamma, self.dbeta, self.mask_y, self.raw_out = dgamma, dbeta, mask_y, raw_out

    def fill(self, f, device):
        from . import nn as K
        f.mode = 2
        f.scratch = K.bn_scratch(device).data_ptr()
        f.gamma = self.gamma.data_ptr()
        f.save_mean, f.save_invstd = self.save_mean.data_ptr(), self.save_invstd.data_ptr()
        f.bn_x, f.mask_y = self.bn_x.data_ptr(), _ptr(self.mask_y) or None
        f.raw_out = _ptr(self.raw_out) or None
        f.dgamma, f.dbeta = self.dgamma.data_ptr(), self.dbeta.data_ptr()

    def cpu(self, g: torch.Tensor, out: torch.Tensor) -> None:
        """g: fp32 GEMM result + addend (not yet masked)."""
        C = g.shape[-1]
        M = g.numel() // C
        g = g.reshape(M, C)
        if self.mask_y is not None:
            g = g * (self.mask_y.float().reshape(M, C) > 0)
        g = g.to(out.dtype).float()
        if self.raw_out is not None:
            self.raw_out.copy_(g.reshape(self.raw_out.shape))
        xhat = (self.bn_x.float().reshape(M, C) - self.save_mean) * self.save_invstd
        s1, s2 = g.sum(0), (g * xhat).sum(0)
        self.dbeta.copy_(s1)
        self.dgamma.copy_(s2)
        out.copy_((self.gamma * self.save_invstd * (g - s1 / M - xhat * s2 / M)).reshape(out.shape))


def max_coresident(block_n: int, a_mn: bool, b_mn: bool, cluster: int) -> int:
    """CTAs of the cluster-mode GEMM variant guaranteed to be resident together on the current GPU (the
    bound a fused-BatchNorm launch, which synchronises its grid in software, must respect)."""
    return int(_native.lib().kml_gemm_max_coresident(int(block_n), int(a_mn), int(b_mn), int(cluster)))


def _epi(bias=None, addend=None, relu=False, out_f32=False, splits=1, block_n=0, slabs=False,
         cluster=False, stats=None, mask_y=None, bn=None, bn_fuse=None, device=None):
    e = _native.Epilogue()
    if bn_fuse is not None:
        f = _native.BnFuse()
        bn_fuse.fill(f, device)
        e._bn_keep = f                        # keep the struct alive for the duration of the call
        import ctypes as _ct
        e.bn_fuse = _ct.pointer(f)
    e.bias = _ptr(bias) or None
    e.addend_bf16 = _ptr(addend) or None
    e.relu, e.out_f32, e.splits, e.block_n = int(relu), int(out_f32), int(splits), int(block_n)
    e.slabs = int(slabs)
    e.cluster = int(cluster)
    e.stats = _ptr(stats) or None
    e.mask_y = _ptr(mask_y) or None
    if bn is not None:                       # (x saved by forward, mean, invstd) of the consuming BatchNorm
        e.bn_x, e.bn_mean, e.bn_invstd = bn[0].data_ptr(), bn[1].data_ptr(), bn[2].data_ptr()
    return e


def _cpu_stats(stats: Optional[torch.Tensor], y: torch.Tensor) -> None:
    """CPU restatement of the cluster epilogue's column statistics: stats[0] += sum, stats[1] += sum of
    squares of the (rounded) output over all rows."""
    if stats is None:
        return
    C = y.shape[-1]
    v = y.float().reshape(-1, C)
    st = stats[: 2 * C].view(2, C)
    st[0] += v.sum(0)
    st[1] += (v * v).sum(0)


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
              bias: Optional[torch.Tensor] = None, relu: bool = False,
              out_dtype: torch.dtype = torch.bfloat16, out: Optional[torch.Tensor] = None,
              splits: int = 1, block_n: int = 0, addend: Optional[torch.Tensor] = None,
              slabs: bool = False, cluster: bool = False, stats: Optional[torch.Tensor] = None,
              bn_fuse=None) -> torch.Tensor:
    """out[M,N] = A · B^T (+bias, ReLU) with fp32 accumulation.

    bn_fuse (`BnFwd` / `BnBwd`, cluster mode, bf16 `out` required): a BatchNorm finished inside this launch.

    cluster=True: the K-splits (<= 8) of a tile run as one thread-block cluster and are added through
    distributed shared memory; the output is final (any dtype, bias/addend/ReLU allowed) and `stats`
    (fp32 [2, N], zeroed) receives per-column sum / sum of squares.

    a: [M,K] (a_mn=False, K contiguous) or [K,M] (a_mn=True, M contiguous), bf16
    b: [N,K] (b_mn=False)               or [K,N] (b_mn=True), bf16
    splits > 1 accumulates K-slices with red.add into an fp32 `out` (zeroed here if not given), or,
    with slabs=True, stores each split's partial result in out[s] (out: fp32 [splits, M, N]).
    """
    M = a.shape[1] if a_mn else a.shape[0]
    K = a.shape[0] if a_mn else a.shape[1]
    N = b.shape[1] if b_mn else b.shape[0]
    if out is not None:
        out_dtype = out.dtype
    if not a.is_cuda:
        A = (a.t() if a_mn else a).float()
        B = (b.t() if b_mn else b).float()
        r = A @ B.t()
        if bias is not None:
            r = r + bias.float()
        if addend is not None:
            r = r + addend.float()
        if relu:
            r = torch.relu(r)
        if bn_fuse is not None:
            bn_fuse.cpu(r.to(out_dtype) if isinstance(bn_fuse, BnFwd) else r, out)
            return out
        r = r.to(out_dtype)
        _cpu_stats(stats, r)
        if out is not None:
            if cluster:
                out.copy_(r)
            elif slabs and splits > 1:
                out.zero_()
                out[0].copy_(r)
            elif splits == 1:
                out.copy_(r)
            else:
                out.add_(r)
            return out
        return r
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16, "gemm_bf16 wants bf16 operands"
    assert a.stride(-1) == 1 and b.stride(-1) == 1 and (out is None or out.stride(-1) == 1)
    out_f32 = out_dtype == torch.float32
    if out is None:
        out = (torch.zeros if splits > 1 and not cluster else torch.empty)((M, N), dtype=out_dtype, device=a.device)
    import ctypes
    e = _epi(bias, addend, relu, out_f32, splits, block_n, slabs and splits > 1 and not cluster, cluster, stats,
             bn_fuse=bn_fuse, device=a.device)
    _native.check(_native.lib().kml_gemm_bf16(
        a.data_ptr(), int(a_mn), a.stride(0), b.data_ptr(), int(b_mn), b.stride(0), out.data_ptr(),
        out.stride(-2), M, N, K, ctypes.byref(e), _native.stream_ptr()), "kml_gemm_bf16")
    return out


def conv2d_fprop_supported(H: int, W: int, Cin: int, Cout: int, KH: int, KW: int, pad: int, stride: int = 1) -> bool:
    """Can the implicit GEMM (shifted, strided TMA boxes; no im2col buffer) take this convolution forward and
    for its weight gradient?  (The data gradient is implicit for stride 1 only.)"""
    Ho, Wo = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    pow2 = lambda v: v > 0 and (v & (v - 1)) == 0
    return (Cin % 64 == 0 and Cout % 8 == 0 and pow2(Ho) and pow2(Wo) and Wo * stride <= 128 and KH * KW <= 64
            and 1 <= stride <= 8)


def conv2d_fprop(x: torch.Tensor, w: torch.Tensor, pad: int, *, bias: Optional[torch.Tensor] = None,
                 relu: bool = False, out_dtype: torch.dtype = torch.bfloat16, splits: int = 1,
                 block_n: int = 0, out: Optional[torch.Tensor] = None, slabs: bool = False,
                 cluster: bool = False, stats: Optional[torch.Tensor] = None, bn_fuse=None,
                 stride: int = 1) -> torch.Tensor:
    """Convolution (any stride the TMA box allows), NHWC.  x: [N,H,W,Cin] bf16; w: [Cout,KH,KW,Cin] bf16 -> [N,Ho,Wo,Cout].
    The data gradient of a stride-1 convolution is this same op applied to dY with the flipped,
    channel-transposed filter and pad' = K-1-pad."""
    N, H, W, Cin = x.shape
    Cout, KH, KW, _ = w.shape
    Ho, Wo = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    if out is not None:
        out_dtype = out.dtype
    if not x.is_cuda:
        y = F.conv2d(x.permute(0, 3, 1, 2).float(), w.permute(0, 3, 1, 2).float(),
                     None if bias is None else bias.float(), padding=pad, stride=stride)
        if relu:
            y = torch.relu(y)
        y = y.permute(0, 2, 3, 1).contiguous().to(out_dtype)
        if bn_fuse is not None:
            bn_fuse.cpu(y, out)
            return out
        _cpu_stats(stats, y)
        if out is not None:
            if slabs and not cluster:      # [S, N, Ho, Wo, C]: everything in slab 0, zeros elsewhere
                out.zero_()
                out[0].copy_(y)
            else:
                out.copy_(y)
            return out
        return y
    assert x.is_contiguous() and w.is_contiguous()
    out_f32 = out_dtype == torch.float32
    import ctypes
    y = out if out is not None else \
        (torch.zeros if splits > 1 and not cluster else torch.empty)((N, Ho, Wo, Cout), dtype=out_dtype, device=x.device)
    e = _epi(bias, None, relu, out_f32, splits, block_n, slabs and not cluster, cluster, stats,
             bn_fuse=bn_fuse, device=x.device)
    _native.check(_native.lib().kml_conv2d_fprop_strided(
        x.data_ptr(), w.data_ptr(), y.data_ptr(), N, H, W, Cin, Cout, KH, KW, pad, int(stride), ctypes.byref(e),
        _native.stream_ptr()), "kml_conv2d_fprop")
    return y


def conv2d_dgrad(dy: torch.Tensor, w: torch.Tensor, pad: int, H: int, W: int, *,
                 addend: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                 block_n: int = 0, splits: int = 1, slabs: bool = False, cluster: bool = False,
                 mask_y: Optional[torch.Tensor] = None, bn=None,
                 stats: Optional[torch.Tensor] = None, bn_fuse=None) -> torch.Tensor:
    """Data gradient of `conv2d_fprop`: dy [N,Ho,Wo,Cout] bf16, w the *forward* filter
    [Cout,KH,KW,Cin] bf16 -> dx [N,H,W,Cin] bf16 (+ addend, the residual branch's gradient).
    Cluster mode can hand the result to a BatchNorm backward ready-made: `mask_y` (that layer's
    output, ReLU mask) is applied, and with `bn` = (x, mean, invstd) `stats` receives
    [sum g, sum g*xhat] per channel (see `ops.nn.bn_bwd_apply`)."""
    N, Ho, Wo, Cout = dy.shape
    _, KH, KW, Cin = w.shape
    if not dy.is_cuda:
        dx = torch.nn.grad.conv2d_input((N, Cin, H, W), w.permute(0, 3, 1, 2).float(),
                                        dy.permute(0, 3, 1, 2).float(), padding=pad).permute(0, 2, 3, 1)
        if addend is not None:
            dx = dx + addend.float()
        if bn_fuse is not None:
            bn_fuse.cpu(dx, out)
            return out
        if mask_y is not None:
            dx = dx * (mask_y.float() > 0)
        if bn is not None:
            gq = dx.to(out.dtype if out is not None else dy.dtype).float().reshape(-1, Cin)
            xh = (bn[0].float().reshape(-1, Cin) - bn[1]) * bn[2]
            st = stats[: 2 * Cin].view(2, Cin)
            st[0] += gq.sum(0)
            st[1] += (gq * xh).sum(0)
        if out is not None:
            if slabs and not cluster:
                out.zero_()
                out[0].copy_(dx)
            else:
                out.copy_(dx.contiguous().to(out.dtype))
            return out
        return dx.contiguous().to(dy.dtype)
    import ctypes
    assert dy.is_contiguous() and w.is_contiguous()
    dx = out if out is not None else torch.empty((N, H, W, Cin), dtype=torch.bfloat16, device=dy.device)
    e = _epi(None, addend, False, dx.dtype == torch.float32, splits, block_n, slabs and not cluster, cluster,
             stats, mask_y, bn, bn_fuse=bn_fuse, device=dy.device)
    _native.check(_native.lib().kml_conv2d_dgrad(
        dy.data_ptr(), w.data_ptr(), dx.data_ptr(), N, H, W, Cin, Cout, KH, KW, pad, ctypes.byref(e),
        _native.stream_ptr()), "kml_conv2d_dgrad")
    return dx


def conv2d_wgrad(x: torch.Tensor, dy: torch.Tensor, KH: int, KW: int, pad: int, *,
                 out: Optional[torch.Tensor] = None, splits: int = 1, block_n: int = 0,
                 stride: int = 1) -> torch.Tensor:
    """Weight gradient of the NHWC convolution: dW[Cout,KH,KW,Cin] fp32.  A caller-provided
    `out` must hold zeros (K-splits are combined with red.add; a single split stores)."""
    N, H, W, Cin = x.shape
    Cout = dy.shape[3]
    if not x.is_cuda:
        g = torch.nn.grad.conv2d_weight(x.permute(0, 3, 1, 2).float(), (Cout, Cin, KH, KW),
                                        dy.permute(0, 3, 1, 2).float(), padding=pad, stride=stride)
        g = g.permute(0, 2, 3, 1).contiguous()
        if out is not None:
            out.add_(g)
            return out
        return g
    assert x.is_contiguous() and dy.is_contiguous()
    if out is None:
        out = torch.zeros((Cout, KH, KW, Cin), dtype=torch.float32, device=x.device)
    _native.check(_native.lib().kml_conv2d_wgrad_strided(
        x.data_ptr(), dy.data_ptr(), out.data_ptr(), N, H, W, Cin, Cout, KH, KW, pad, int(stride), int(splits),
        int(block_n), _native.stream_ptr()), "kml_conv2d_wgrad")
    return out
