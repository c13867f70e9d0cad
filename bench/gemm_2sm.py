"""Two-SM GEMM kernel (tcgen05.mma.cta_group::2): correctness against fp32 torch, then timing against the
one-CTA kernels and cuBLAS."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import ops
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)


def timeit(fn, n=10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fn(); torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


torch.manual_seed(0)
for (M, N, K) in [(256, 256, 64), (512, 512, 256), (1024, 768, 4096), (640, 304, 200), (128, 256, 128)]:
    a = torch.randn(M, K, device=dev).bfloat16()
    b = (torch.randn(N, K, device=dev) * 0.1).bfloat16()
    bias = torch.randn(N, device=dev)
    ref = torch.relu(a.float() @ b.float().t() + bias)
    out = torch.full((M, N), 7.0, device=dev, dtype=torch.bfloat16)
    ops.gemm_bf16(a, b, bias=bias, relu=True, out=out, block_n=512)
    torch.cuda.synchronize()
    err = float((out.float() - ref).abs().max() / ref.abs().max())
    print(f"2sm {M}x{N}x{K}: rel err {err:.4f}", flush=True)
    assert err < 2e-2
    of = torch.empty(M, N, device=dev)
    ops.gemm_bf16(a, b, out=of, block_n=512)
    torch.cuda.synchronize()
    e2 = float((of - a.float() @ b.float().t()).abs().max())
    assert e2 < 1e-2 * max(1.0, float(ref.abs().max())), e2
for (M, N, K) in [(8192, 4096, 4096), (16384, 8192, 8192), (4096, 4096, 16384)]:
    a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(N, K, device=dev).bfloat16()
    o = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    fl = 2.0 * M * N * K
    for bn in (128, 256, 512):
        t = timeit(lambda: ops.gemm_bf16(a, b, out=o, block_n=bn))
        print(f"{M}x{N}x{K} block_n {bn}: {t:8.1f} us  {fl / t / 1e6:7.0f} TFLOP/s", flush=True)
    t = timeit(lambda: torch.matmul(a, b.t(), out=o))
    print(f"{M}x{N}x{K} cuBLAS      : {t:8.1f} us  {fl / t / 1e6:7.0f} TFLOP/s", flush=True)
print("ok")
