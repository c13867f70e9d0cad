"""Python entry points of the hand-written sm_100a kernels that are not GEMM-shaped:
flat optimizers, casts, fused softmax-cross-entropy (csrc/optim.cu, csrc/loss.cu).

Dispatch rule used by every op in this package: CUDA tensors go to the native kernel and raise if
the library is missing (no silent PyTorch fallback on a GPU box); CPU tensors take a plain fp32
PyTorch restatement of the same math, which is also what the numerics tests compare against.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from .. import _native


def _ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


# ------------------------------------------------------------------------------------------------
# optimizers over flat buffers
# ------------------------------------------------------------------------------------------------
def sgd_step_(w: torch.Tensor, g: torch.Tensor, momentum_buf: Optional[torch.Tensor], lr: float,
              weight_decay: float = 0.0, momentum: float = 0.0, nesterov: bool = False,
              w_bf16: Optional[torch.Tensor] = None) -> None:
    """In-place torch.optim.SGD step over flat fp32 buffers (a zeroed momentum buffer reproduces
    torch's first-step `buf = grad`)."""
    if w.is_cuda:
        lib = _native.lib()
        _native.check(lib.kml_sgd_step(w.data_ptr(), g.data_ptr(), _ptr(momentum_buf), w.numel(),
                                       lr, weight_decay, momentum, int(nesterov), _ptr(w_bf16),
                                       _native.stream_ptr()), "kml_sgd_step")
        return
    d = g + weight_decay * w if weight_decay else g.clone()
    if momentum:
        momentum_buf.mul_(momentum).add_(d)
        d = d + momentum * momentum_buf if nesterov else momentum_buf
    w.add_(d, alpha=-lr)
    if w_bf16 is not None:
        w_bf16.copy_(w)


def adam_step_(w: torch.Tensor, g: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor,
               step: int, lr: float, betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8,
               weight_decay: float = 0.0, w_bf16: Optional[torch.Tensor] = None,
               step_dev: Optional[torch.Tensor] = None) -> None:
    """`step_dev` (int32[2] on the device, zero at reset) replaces the host `step`: the kernel keeps
    the count itself, which is what a CUDA-graph replay needs."""
    if w.is_cuda and step_dev is not None:
        _native.check(_native.lib().kml_adam_step_dev(
            w.data_ptr(), g.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(), w.numel(), lr, betas[0], betas[1],
            eps, weight_decay, step_dev.data_ptr(), _ptr(w_bf16), _native.stream_ptr()), "kml_adam_step_dev")
        return
    if w.is_cuda:
        lib = _native.lib()
        _native.check(lib.kml_adam_step(w.data_ptr(), g.data_ptr(), exp_avg.data_ptr(),
                                        exp_avg_sq.data_ptr(), w.numel(), lr, betas[0], betas[1],
                                        eps, weight_decay, int(step), _ptr(w_bf16),
                                        _native.stream_ptr()), "kml_adam_step")
        return
    d = g + weight_decay * w if weight_decay else g
    exp_avg.mul_(betas[0]).add_(d, alpha=1 - betas[0])
    exp_avg_sq.mul_(betas[1]).addcmul_(d, d, value=1 - betas[1])
    bc1 = 1 - betas[0] ** step
    bc2 = 1 - betas[1] ** step
    denom = exp_avg_sq.sqrt() / (bc2 ** 0.5) + eps
    w.addcdiv_(exp_avg, denom, value=-lr / bc1)
    if w_bf16 is not None:
        w_bf16.copy_(w)


def fill_(t: torch.Tensor, value: float) -> None:
    if t.is_cuda and t.dtype == torch.float32 and t.numel() % 4 == 0 and t.data_ptr() % 16 == 0:
        _native.check(_native.lib().kml_fill_f32(t.data_ptr(), float(value), t.numel(),
                                                 _native.stream_ptr()), "kml_fill_f32")
    else:
        t.fill_(value)


def cast_bf16_(dst: torch.Tensor, src: torch.Tensor) -> None:
    if src.is_cuda:
        _native.check(_native.lib().kml_cast_f32_to_bf16(src.data_ptr(), dst.data_ptr(), src.numel(),
                                                         _native.stream_ptr()), "kml_cast_f32_to_bf16")
    else:
        dst.copy_(src)


# ------------------------------------------------------------------------------------------------
# fused softmax cross-entropy
# ------------------------------------------------------------------------------------------------
_ce_scratch = {}


def _scratch_for(device: torch.device, rows: int) -> torch.Tensor:
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    need = int(_native.lib().kml_softmax_ce_scratch_bytes(rows))
    buf = _ce_scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.zeros(max(need, 4096), dtype=torch.uint8, device=device)
        _ce_scratch[key] = buf
    return buf


def softmax_ce_raw(logits: torch.Tensor, labels: torch.Tensor, want_grad: bool = True,
                   want_correct: bool = False):
    """(loss[1] fp32, dlogits or None, correct[1] int32 or None); dlogits is d(mean loss)/dlogits."""
    rows, classes = logits.shape
    if logits.is_cuda:
        logits = logits.contiguous()
        labels = labels.contiguous()
        loss = torch.empty(1, dtype=torch.float32, device=logits.device)
        dl = torch.empty_like(logits) if want_grad else None
        corr = torch.empty(1, dtype=torch.int32, device=logits.device) if want_correct else None
        is_bf16 = logits.dtype == torch.bfloat16
        if not is_bf16 and logits.dtype != torch.float32:
            raise TypeError("softmax_ce: logits must be fp32 or bf16")
        _native.check(_native.lib().kml_softmax_ce(
            logits.data_ptr(), int(is_bf16), labels.data_ptr(), rows, classes, loss.data_ptr(),
            _ptr(dl), _ptr(corr), _scratch_for(logits.device, rows).data_ptr(),
            _native.stream_ptr()), "kml_softmax_ce")
        return loss, dl, corr
    x = logits.float()
    lse = torch.logsumexp(x, dim=1)
    loss = (lse - x.gather(1, labels.view(-1, 1)).squeeze(1)).mean().view(1)
    dl = None
    if want_grad:
        dl = torch.softmax(x, dim=1)
        dl[torch.arange(rows), labels] -= 1.0
        dl = (dl / rows).to(logits.dtype)
    corr = x.argmax(dim=1).eq(labels).sum().to(torch.int32).view(1) if want_correct else None
    return loss, dl, corr


class _SoftmaxCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        loss, dl, _ = softmax_ce_raw(logits, labels, want_grad=True)
        ctx.save_for_backward(dl)
        return loss.view(())

    @staticmethod
    def backward(ctx, grad_out):
        (dl,) = ctx.saved_tensors
        return dl * grad_out.to(dl.dtype), None


def cross_entropy(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """Drop-in for F.cross_entropy(logits, labels) (mean reduction): forward and backward are
    produced by one kernel launch."""
    return _SoftmaxCE.apply(logits, labels)


def accuracy_count(logits: torch.Tensor, labels: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """(loss, number of correct predictions) with no host sync."""
    loss, _, corr = softmax_ce_raw(logits, labels, want_grad=False, want_correct=True)
    return loss.view(()), corr.view(())
