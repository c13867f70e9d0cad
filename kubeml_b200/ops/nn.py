"""Python entry points of the NHWC bf16 layer kernels (csrc/norm.cu, csrc/nn_misc.cu).  Raw,
stateless calls on preallocated tensors — the native models (kubeml_b200.models) own the buffers
and the order of calls; CPU tensors take an fp32 PyTorch restatement (test oracle)."""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from .. import _native


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


_bn_scratch = {}


def bn_scratch(device: torch.device, C: int = 0) -> torch.Tensor:
    """Launch state of the BatchNorm kernels (double-buffered statistics + launch counter): zeroed
    once here, then owned by the kernels.  One per (device, stream): launches on a stream serialise."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    buf = _bn_scratch.get(key)
    if buf is None:
        buf = torch.zeros(int(_native.lib().kml_bn_scratch_bytes(C)), dtype=torch.uint8, device=device)
        _bn_scratch[key] = buf
    return buf


def _fold(x, x_slabs):
    """CPU restatement: the layer input is x, or the sum of the split-K slabs."""
    return x.float() if x_slabs is None else x_slabs.float().sum(0)


def bn_fwd_train(x, res, y, gamma, beta, running_mean, running_var, nbt, save_mean, save_invstd,
                 relu: bool, eps: float = 1e-5, momentum: float = 0.1, x_slabs=None, x_out=None) -> None:
    """y = relu?(bn(x) + res?) with batch statistics.  Input is `x` (bf16 [..., C]) or `x_slabs`
    (fp32 [S, ..., C], split-K partials of the producing GEMM); in the slab case `x_out` receives the
    folded bf16 input that backward needs."""
    C = y.shape[-1]
    M = y.numel() // C
    if y.is_cuda:
        S = 0 if x_slabs is None else x_slabs.shape[0]
        _native.check(_native.lib().kml_bn_fwd_train(
            _ptr(x), _ptr(x_slabs), S, 0 if x_slabs is None else x_slabs.stride(0), _ptr(x_out), _ptr(res),
            y.data_ptr(), gamma.data_ptr(), beta.data_ptr(), _ptr(running_mean), _ptr(running_var), _ptr(nbt),
            save_mean.data_ptr(), save_invstd.data_ptr(), bn_scratch(y.device, C).data_ptr(), M, C, int(relu),
            eps, momentum, _native.stream_ptr()), "kml_bn_fwd_train")
        return
    xf = _fold(x, x_slabs).reshape(M, C)
    if x_out is not None:
        x_out.copy_(xf.reshape(x_out.shape))
        xf = x_out.float().reshape(M, C)
    mean = xf.mean(0)
    var = xf.var(0, unbiased=False)
    invstd = torch.rsqrt(var + eps)
    out = (xf - mean) * invstd * gamma + beta
    if res is not None:
        out = out + res.float().reshape(M, C)
    if relu:
        out = torch.relu(out)
    y.copy_(out.reshape(y.shape))
    save_mean.copy_(mean)
    save_invstd.copy_(invstd)
    if running_mean is not None:
        running_mean.mul_(1 - momentum).add_(momentum * mean)
        running_var.mul_(1 - momentum).add_(momentum * var * (M / max(M - 1, 1)))
    if nbt is not None:
        nbt.add_(1)


def bn_stats_numel(C: int) -> int:
    """fp32 elements of a statistics buffer for `bn_apply_train` ([2][C] sums + ticket word)."""
    return 2 * C + 4


def bn_apply_train(x, stats, res, y, gamma, beta, running_mean, running_var, nbt, save_mean, save_invstd,
                   relu: bool, eps: float = 1e-5, momentum: float = 0.1) -> None:
    """BatchNorm forward from statistics the producing GEMM accumulated (`stats`: fp32 [2C+4], sums and
    sums of squares over the M rows; cleared again here)."""
    C = y.shape[-1]
    M = y.numel() // C
    if y.is_cuda:
        _native.check(_native.lib().kml_bn_apply_train(
            x.data_ptr(), stats.data_ptr(), _ptr(res), y.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
            _ptr(running_mean), _ptr(running_var), _ptr(nbt), save_mean.data_ptr(), save_invstd.data_ptr(), M, C,
            int(relu), eps, momentum, _native.stream_ptr()), "kml_bn_apply_train")
        return
    st = stats[: 2 * C].view(2, C)
    mean = st[0] / M
    var = (st[1] / M - mean * mean).clamp_min(0.0)
    invstd = torch.rsqrt(var + eps)
    out = (x.float().reshape(M, C) - mean) * invstd * gamma + beta
    if res is not None:
        out = out + res.float().reshape(M, C)
    if relu:
        out = torch.relu(out)
    y.copy_(out.reshape(y.shape))
    save_mean.copy_(mean)
    save_invstd.copy_(invstd)
    if running_mean is not None:
        running_mean.mul_(1 - momentum).add_(momentum * mean)
        running_var.mul_(1 - momentum).add_(momentum * var * (M / max(M - 1, 1)))
    if nbt is not None:
        nbt.add_(1)
    stats.zero_()


def bn_bwd_apply(g, x, gamma, save_mean, save_invstd, stats, dx, dgamma, dbeta) -> None:
    """BatchNorm backward from the masked gradient `g` and the sums [sum g, sum g*xhat] the producing
    data-gradient GEMM accumulated in `stats` (cleared again here)."""
    C = x.shape[-1]
    M = x.numel() // C
    if x.is_cuda:
        _native.check(_native.lib().kml_bn_bwd_apply(
            g.data_ptr(), x.data_ptr(), gamma.data_ptr(), save_mean.data_ptr(), save_invstd.data_ptr(),
            stats.data_ptr(), dx.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), M, C, _native.stream_ptr()),
            "kml_bn_bwd_apply")
        return
    st = stats[: 2 * C].view(2, C)
    s1, s2 = st[0].clone(), st[1].clone()
    gf = g.float().reshape(M, C)
    xhat = (x.float().reshape(M, C) - save_mean) * save_invstd
    dbeta.copy_(s1)
    dgamma.copy_(s2)
    dx.copy_((gamma * save_invstd * (gf - s1 / M - xhat * s2 / M)).reshape(dx.shape))
    stats.zero_()


def bn_fwd_eval(x, res, y, gamma, beta, running_mean, running_var, relu: bool, eps: float = 1e-5,
                x_slabs=None) -> None:
    C = y.shape[-1]
    M = y.numel() // C
    if y.is_cuda:
        S = 0 if x_slabs is None else x_slabs.shape[0]
        _native.check(_native.lib().kml_bn_fwd_eval(
            _ptr(x), _ptr(x_slabs), S, 0 if x_slabs is None else x_slabs.stride(0), _ptr(res), y.data_ptr(),
            gamma.data_ptr(), beta.data_ptr(), running_mean.data_ptr(), running_var.data_ptr(), M, C, int(relu),
            eps, _native.stream_ptr()), "kml_bn_fwd_eval")
        return
    out = (_fold(x, x_slabs).reshape(M, C) - running_mean) * torch.rsqrt(running_var + eps) * gamma + beta
    if res is not None:
        out = out + res.float().reshape(M, C)
    if relu:
        out = torch.relu(out)
    y.copy_(out.reshape(y.shape))


def bn_bwd(dy, y, x, gamma, save_mean, save_invstd, dx, dres, dgamma, dbeta, dy_slabs=None, dy_add=None) -> None:
    """dy (bf16) or dy_slabs (fp32 [S, ...]) [+ dy_add]: grad of the block output; y: block output
    for the ReLU mask (None = no ReLU)."""
    C = x.shape[-1]
    M = x.numel() // C
    if x.is_cuda:
        S = 0 if dy_slabs is None else dy_slabs.shape[0]
        _native.check(_native.lib().kml_bn_bwd(
            _ptr(dy), _ptr(dy_slabs), S, 0 if dy_slabs is None else dy_slabs.stride(0), _ptr(dy_add), _ptr(y),
            x.data_ptr(), gamma.data_ptr(), save_mean.data_ptr(), save_invstd.data_ptr(), dx.data_ptr(), _ptr(dres),
            dgamma.data_ptr(), dbeta.data_ptr(), bn_scratch(x.device, C).data_ptr(), M, C, _native.stream_ptr()),
            "kml_bn_bwd")
        return
    g = _fold(dy, dy_slabs).reshape(M, C)
    if dy_add is not None:
        g = g + dy_add.float().reshape(M, C)
    if y is not None:
        g = g * (y.float().reshape(M, C) > 0)
    xhat = (x.float().reshape(M, C) - save_mean) * save_invstd
    s1, s2 = g.sum(0), (g * xhat).sum(0)
    dbeta.copy_(s1)
    dgamma.copy_(s2)
    if dres is not None:
        dres.copy_(g.reshape(dres.shape))
    dx.copy_((gamma * save_invstd * (g - s1 / M - xhat * s2 / M)).reshape(dx.shape))


def maxpool_fwd(x, y, idx, k: int, stride: int, pad: int) -> None:
    N, H, W, C = x.shape
    if x.is_cuda:
        _native.check(_native.lib().kml_maxpool_fwd(x.data_ptr(), y.data_ptr(), _ptr(idx), N, H, W, C, k, stride,
                                                    pad, _native.stream_ptr()), "kml_maxpool_fwd")
        return
    out, ind = F.max_pool2d(x.float().permute(0, 3, 1, 2), k, stride, pad, return_indices=True)
    y.copy_(out.permute(0, 2, 3, 1))
    if idx is not None:       # convert flat input index -> tap index within the window
        Ho, Wo = out.shape[2], out.shape[3]
        ih, iw = ind // W, ind % W
        ho = torch.arange(Ho).view(1, 1, Ho, 1)
        wo = torch.arange(Wo).view(1, 1, 1, Wo)
        tap = (ih - (ho * stride - pad)) * k + (iw - (wo * stride - pad))
        idx.copy_(tap.permute(0, 2, 3, 1).to(torch.uint8))


def maxpool_bwd(dy, idx, dx, k: int, stride: int, pad: int) -> None:
    N, H, W, C = dx.shape
    if dy.is_cuda:
        _native.check(_native.lib().kml_maxpool_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, C, k,
                                                    stride, pad, _native.stream_ptr()), "kml_maxpool_bwd")
        return
    Ho, Wo = dy.shape[1], dy.shape[2]
    out = torch.zeros(N, H, W, C)
    t = idx.long()
    kh, kw = t // k, t % k
    ho = torch.arange(Ho).view(1, Ho, 1, 1)
    wo = torch.arange(Wo).view(1, 1, Wo, 1)
    h = ho * stride - pad + kh
    w = wo * stride - pad + kw
    n = torch.arange(N).view(N, 1, 1, 1).expand_as(h)
    c = torch.arange(C).view(1, 1, 1, C).expand_as(h)
    out.index_put_((n, h, w, c), dy.float(), accumulate=True)
    dx.copy_(out)


def avgpool_fwd(x, y) -> None:
    N, H, W, C = x.shape
    if x.is_cuda:
        _native.check(_native.lib().kml_avgpool_fwd(x.data_ptr(), y.data_ptr(), N, H * W, C,
                                                    _native.stream_ptr()), "kml_avgpool_fwd")
        return
    y.copy_(x.float().mean(dim=(1, 2)))


def avgpool_bwd(dy, dx) -> None:
    N, H, W, C = dx.shape
    if dy.is_cuda:
        _native.check(_native.lib().kml_avgpool_bwd(dy.data_ptr(), dx.data_ptr(), N, H * W, C,
                                                    _native.stream_ptr()), "kml_avgpool_bwd")
        return
    dx.copy_((dy.float() / (H * W)).view(N, 1, 1, C).expand(N, H, W, C))


def im2col(x, cols, KH: int, KW: int, stride: int, pad: int) -> None:
    """cols: bf16 [M, ldk] with ldk >= KH*KW*C; column order (kh, kw, c)."""
    N, H, W, C = x.shape
    ldk = cols.shape[1]
    if x.is_cuda:
        _native.check(_native.lib().kml_im2col(x.data_ptr(), cols.data_ptr(), N, H, W, C, KH, KW, stride, pad, ldk,
                                               _native.stream_ptr()), "kml_im2col")
        return
    u = F.unfold(x.float().permute(0, 3, 1, 2), (KH, KW), padding=pad, stride=stride)   # [N, C*KH*KW, L]
    L = u.shape[2]
    u = u.view(N, C, KH * KW, L).permute(0, 3, 2, 1).reshape(N * L, KH * KW * C)
    cols.zero_()
    cols[:, :KH * KW * C].copy_(u)


def col2im(dcols, addend, dx, KH: int, KW: int, stride: int, pad: int) -> None:
    N, H, W, C = dx.shape
    ldk = dcols.shape[1]
    if dcols.is_cuda:
        _native.check(_native.lib().kml_col2im(dcols.data_ptr(), _ptr(addend), dx.data_ptr(), N, H, W, C, KH, KW,
                                               stride, pad, ldk, _native.stream_ptr()), "kml_col2im")
        return
    Ho, Wo = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    u = dcols[:, :KH * KW * C].float().view(N, Ho * Wo, KH * KW, C).permute(0, 3, 2, 1).reshape(N, C * KH * KW, Ho * Wo)
    out = F.fold(u, (H, W), (KH, KW), padding=pad, stride=stride).permute(0, 2, 3, 1)
    if addend is not None:
        out = out + addend.float()
    dx.copy_(out)


def colsum(x, out) -> None:
    if x.is_cuda:
        _native.check(_native.lib().kml_colsum_bf16(x.data_ptr(), out.data_ptr(), x.shape[0], x.shape[1], x.stride(0),
                                                    _native.stream_ptr()), "kml_colsum_bf16")
        return
    out.copy_(x.float().sum(0))


def stamp(slots, idx: int) -> None:
    """slots[idx] (int64 on the device) = the GPU's nanosecond timer when the stream gets here."""
    if slots.is_cuda:
        _native.check(_native.lib().kml_stamp(slots.data_ptr(), int(idx), _native.stream_ptr()), "kml_stamp")


def post_f32(src, dst_pinned) -> None:
    """dst_pinned (page-locked host float32, same number of elements) <- src (device float32), written by a kernel
    on the current stream; read it after synchronising with that stream."""
    assert src.is_cuda and src.dtype == torch.float32 and dst_pinned.dtype == torch.float32
    assert dst_pinned.is_pinned() and src.is_contiguous() and dst_pinned.is_contiguous()
    _native.check(_native.lib().kml_post_f32(src.data_ptr(), dst_pinned.data_ptr(), src.numel(), _native.stream_ptr()),
                  "kml_post_f32")


def pack_rows(src, dst) -> None:
    """dst bf16 [rows, ld] <- src fp32 [rows, cols], zero padded."""
    rows, cols = src.shape
    if src.is_cuda:
        _native.check(_native.lib().kml_pack_rows_bf16(src.data_ptr(), dst.data_ptr(), rows, cols, dst.shape[1],
                                                       _native.stream_ptr()), "kml_pack_rows_bf16")
        return
    dst.zero_()
    dst[:, :cols].copy_(src)


def image_prep(x_u8, y, aug, pad: int, mean: Tuple[float, float, float], std: Tuple[float, float, float],
               rand_state=None, seed: int = 0, labels=None, labels_out=None) -> None:
    """uint8 NHWC -> normalised bf16 NHWC with optional per-sample flip/crop codes (see header).  With
    `rand_state` (int64[2] on the device, zeros) the codes are drawn inside the kernel instead of being passed;
    that kernel can also carry the batch's int64 `labels` into `labels_out` (a fixed buffer) in the same launch."""
    N, H, W, C = x_u8.shape
    if x_u8.is_cuda and rand_state is not None:
        m = (ctypes.c_float * 3)(*mean)
        s = (ctypes.c_float * 3)(*std)
        if labels_out is not None:
            assert labels.dtype == torch.int64 and labels_out.dtype == torch.int64 and labels.numel() == N
            assert labels.is_contiguous() and labels_out.is_contiguous() and labels_out.numel() >= N
        _native.check(_native.lib().kml_image_prep_rand(x_u8.data_ptr(), y.data_ptr(), rand_state.data_ptr(),
                                                        int(seed) & 0xFFFFFFFF, N, H, W, C, pad, m, s,
                                                        labels.data_ptr() if labels_out is not None else None,
                                                        labels_out.data_ptr() if labels_out is not None else None,
                                                        _native.stream_ptr()), "kml_image_prep_rand")
        return
    if labels_out is not None:
        labels_out[:N].copy_(labels)
    if x_u8.is_cuda:
        m = (ctypes.c_float * 3)(*mean)
        s = (ctypes.c_float * 3)(*std)
        _native.check(_native.lib().kml_image_prep(x_u8.data_ptr(), y.data_ptr(), _ptr(aug), N, H, W, C, pad, m, s,
                                                   _native.stream_ptr()), "kml_image_prep")
        return
    xf = x_u8.float() / 255.0
    if aug is not None:
        out = torch.zeros_like(xf)
        for n in range(N):
            a = int(aug[n])
            dy, dxo = ((a >> 8) & 0xff) - pad, ((a >> 16) & 0xff) - pad
            img = xf[n].flip(1) if a & 1 else xf[n]
            padded = F.pad(img.permute(2, 0, 1), (pad, pad, pad, pad)).permute(1, 2, 0)
            out[n] = padded[pad + dy: pad + dy + H, pad + dxo: pad + dxo + W]
        xf = out
    y.copy_((xf - torch.tensor(mean)) / torch.tensor(std))


def relu_bwd(dy, y, g, scale: float = 1.0) -> None:
    """g = scale * dy * (y > 0)."""
    if dy.is_cuda:
        _native.check(_native.lib().kml_relu_bwd(dy.data_ptr(), y.data_ptr(), g.data_ptr(), scale, dy.numel(),
                                                 _native.stream_ptr()), "kml_relu_bwd")
        return
    g.copy_(dy.float() * (y.float() > 0) * scale)


def dropout_fwd(x, y, mask, counter, layer: int, p: float) -> None:
    """Inverted dropout; `counter` is a device int64 the caller bumps once per step (`bump_counter`)."""
    if x.is_cuda:
        _native.check(_native.lib().kml_dropout_fwd(x.data_ptr(), y.data_ptr(), _ptr(mask), counter.data_ptr(), layer,
                                                    p, x.numel(), _native.stream_ptr()), "kml_dropout_fwd")
        return
    keep = torch.rand(x.shape) >= p
    if mask is not None:
        mask.copy_(keep.to(torch.uint8))
    y.copy_(x.float() * keep / (1.0 - p))


def dropout_bwd(dy, mask, dx, p: float) -> None:
    if dy.is_cuda:
        _native.check(_native.lib().kml_dropout_bwd(dy.data_ptr(), mask.data_ptr(), dx.data_ptr(), p, dy.numel(),
                                                    _native.stream_ptr()), "kml_dropout_bwd")
        return
    dx.copy_(dy.float() * mask.float() / (1.0 - p))


def bump_counter(counter) -> None:
    if counter.is_cuda:
        _native.check(_native.lib().kml_bump_counter(counter.data_ptr(), _native.stream_ptr()), "kml_bump_counter")
        return
    counter.add_(1)


def replicate_fwd(x, y, r2: int) -> None:
    """y[n, c*r2 + r] = x[n, c]: AdaptiveAvgPool2d((R,R)) + flatten of a 1x1 feature map."""
    N, C = x.shape
    if x.is_cuda:
        _native.check(_native.lib().kml_replicate_fwd(x.data_ptr(), y.data_ptr(), N, C, r2, _native.stream_ptr()),
                      "kml_replicate_fwd")
        return
    y.copy_(x.unsqueeze(2).expand(N, C, r2).reshape(N, C * r2))


def replicate_bwd(dy, dx, r2: int) -> None:
    N, C = dx.shape
    if dy.is_cuda:
        _native.check(_native.lib().kml_replicate_bwd(dy.data_ptr(), dx.data_ptr(), N, C, r2, _native.stream_ptr()),
                      "kml_replicate_bwd")
        return
    dx.copy_(dy.float().view(N, C, r2).sum(2))


def relu_bwd_cols(dy, y, g, scale: float = 1.0, dbias=None, dy_slabs=None) -> None:
    """g = scale * v * (y > 0), v = dy (bf16 [M, C]) or the fold of fp32 split-K slabs [S, M, C];
    `dbias` (fp32 [C]) receives the column sums of g in the same pass."""
    C = y.shape[-1]
    M = y.numel() // C
    if y.is_cuda:
        S = 0 if dy_slabs is None else dy_slabs.shape[0]
        _native.check(_native.lib().kml_relu_bwd_cols(
            _ptr(dy), _ptr(dy_slabs), S, 0 if dy_slabs is None else dy_slabs.stride(0), y.data_ptr(), g.data_ptr(),
            scale, _ptr(dbias), M, C, _native.stream_ptr()), "kml_relu_bwd_cols")
        return
    v = _fold(dy, dy_slabs).reshape(M, C) * (y.float().reshape(M, C) > 0) * scale
    g.copy_(v.reshape(g.shape))
    if dbias is not None:
        dbias.copy_(v.sum(0))


def bias_act(x_slabs, bias, y, relu: bool) -> None:
    """y (bf16 [M, C], row stride may be padded) = relu?(sum_s x_slabs[s] + bias)."""
    S = x_slabs.shape[0]
    C = y.shape[-1]
    M = y.numel() // C if y.is_contiguous() else y.shape[0]
    if y.is_cuda:
        xs = x_slabs.view(S, M, -1)
        _native.check(_native.lib().kml_bias_act(
            xs.data_ptr(), S, xs.stride(0), _ptr(bias), y.data_ptr(), int(relu), M, C, xs.stride(1),
            y.stride(0) if y.dim() == 2 else C, _native.stream_ptr()), "kml_bias_act")
        return
    out = x_slabs.float().sum(0).reshape(M, -1)[:, :C]
    if bias is not None:
        out = out + bias
    if relu:
        out = torch.relu(out)
    y.copy_(out.reshape(y.shape))
