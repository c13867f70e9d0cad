"""Diagnostic sweep of the tcgen05 GEMM family: prints the error of every variant/shape instead of
stopping at the first failure (used while bringing the kernels up; `pytest -m gpu` is the gate)."""
import sys, os, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import ops

dev = torch.device("cuda", 0)
torch.manual_seed(0)


def rel(got, want):
    return float((got.float() - want).abs().max() / want.abs().max().clamp_min(1e-6))


def gemm_case(M, N, K, a_mn, b_mn, block_n=0, splits=1):
    A = torch.randn(M, K, device=dev).bfloat16()
    B = torch.randn(N, K, device=dev).bfloat16()
    want = A.float() @ B.float().t()
    a = A.t().contiguous() if a_mn else A
    b = B.t().contiguous() if b_mn else B
    got = ops.gemm_bf16(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32, block_n=block_n, splits=splits)
    torch.cuda.synchronize()
    e = rel(got, want)
    tag = "ok " if e < 2e-3 else "BAD"
    extra = ""
    if e >= 2e-3:
        # where is it wrong?  per 32x32 block error map summary
        d = (got - want).abs()
        rows_bad = (d.max(dim=1).values > 1e-2 * want.abs().max()).nonzero().flatten()
        cols_bad = (d.max(dim=0).values > 1e-2 * want.abs().max()).nonzero().flatten()
        extra = f" bad_rows[{rows_bad.numel()}]={rows_bad[:6].tolist()} bad_cols[{cols_bad.numel()}]={cols_bad[:6].tolist()} got00={got[0,0].item():.3f} want00={want[0,0].item():.3f}"
    print(f"{tag} gemm M{M} N{N} K{K} a_mn={int(a_mn)} b_mn={int(b_mn)} bn={block_n} s={splits} rel={e:.2e}{extra}", flush=True)


def conv_case(N, H, C, Co, k, pad):
    x = torch.randn(N, H, H, C, device=dev).bfloat16()
    w = (torch.randn(Co, k, k, C, device=dev) / (k * k * C) ** 0.5).bfloat16()
    want = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), padding=pad).permute(0, 2, 3, 1)
    got = ops.conv2d_fprop(x, w, pad, out_dtype=torch.float32)
    torch.cuda.synchronize()
    e = rel(got, want)
    print(f"{'ok ' if e < 2e-3 else 'BAD'} fprop N{N} H{H} C{C} Co{Co} k{k} rel={e:.2e}", flush=True)
    Ho = H + 2 * pad - k + 1
    dy = torch.randn(N, Ho, Ho, Co, device=dev).bfloat16()
    wantg = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (Co, C, k, k), dy.float().permute(0, 3, 1, 2), padding=pad).permute(0, 2, 3, 1)
    gotg = ops.conv2d_wgrad(x, dy, k, k, pad)
    torch.cuda.synchronize()
    e = rel(gotg, wantg)
    print(f"{'ok ' if e < 2e-3 else 'BAD'} wgrad N{N} H{H} C{C} Co{Co} k{k} rel={e:.2e}", flush=True)


cases = []
for amn, bmn in ((False, False), (False, True), (True, True)):
    for (M, N, K) in ((128, 64, 64), (128, 64, 128), (128, 128, 64), (256, 256, 512), (128, 64, 16)):
        cases.append((gemm_case, (M, N, K, amn, bmn)))
cases.append((gemm_case, (256, 128, 1024, False, False, 64, 4)))
for c in ((2, 8, 64, 64, 3, 1), (8, 4, 128, 128, 3, 1), (128, 1, 512, 512, 3, 1), (2, 16, 64, 64, 3, 1), (4, 8, 64, 64, 1, 0)):
    cases.append((conv_case, c))
for fn, a in cases:
    try:
        fn(*a)
    except Exception as e:
        print("EXC", fn.__name__, a, repr(e)[:300], flush=True)
        if "CUDA" in repr(e) or "cuda" in repr(e):
            print("aborting sweep after a CUDA error (context is poisoned)")
            break
