"""Implicit-GEMM convolution (shifted 4-D TMA boxes) against the plain GEMM of the same M x N x K:
how much does the convolution addressing cost?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import ops
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)


def timeit(fn, n=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fn(); torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


for (N, H, C, Co) in [(128, 32, 64, 64), (128, 16, 128, 128), (128, 8, 256, 256), (128, 4, 512, 512), (128, 2, 512, 512)]:
    x = torch.randn(N, H, H, C, device=dev).bfloat16()
    w = (torch.randn(Co, 3, 3, C, device=dev) * 0.05).bfloat16()
    y = torch.empty(N, H, H, Co, device=dev, dtype=torch.bfloat16)
    M, K = N * H * H, 9 * C
    A = torch.randn(M, K, device=dev).bfloat16()
    Bm = w.view(Co, K)
    o = torch.empty(M, Co, device=dev, dtype=torch.bfloat16)
    fl = 2.0 * M * Co * K
    for bn in (64, 128):
        if bn > Co: continue
        tc = timeit(lambda: ops.conv2d_fprop(x, w, 1, out=y, block_n=bn))
        tg = timeit(lambda: ops.gemm_bf16(A, Bm, out=o, block_n=bn))
        print(f"N{N} {H}x{H} C{C}->{Co} block_n {bn}: conv {tc:7.1f} us ({fl / tc / 1e6:6.0f} TFLOP/s)   gemm {tg:7.1f} us ({fl / tg / 1e6:6.0f} TFLOP/s)")
    xc = x.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
    wc = w.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
    t = timeit(lambda: torch.nn.functional.conv2d(xc, wc, padding=1))
    print(f"   cuDNN bf16 channels_last: {t:7.1f} us ({fl / t / 1e6:6.0f} TFLOP/s)")
