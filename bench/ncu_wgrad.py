"""One launch of the VGG fc1 weight-gradient GEMM (dW[4096,25088] = dY^T[4096,128] . X[128,25088], fp32 out)
and of the big square GEMM, for `ncu --set full`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import ops
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
dy = torch.randn(128, 4096, device=dev).bfloat16()
x = torch.randn(128, 25088, device=dev).bfloat16()
dw = torch.empty(4096, 25088, device=dev)
for _ in range(3):
    ops.gemm_bf16(dy, x, a_mn=True, b_mn=True, out_dtype=torch.float32, out=dw)
torch.cuda.synchronize()
def timeit(fn, n=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fn(); torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n
print("fc1 wgrad us", timeit(lambda: ops.gemm_bf16(dy, x, a_mn=True, b_mn=True, out_dtype=torch.float32, out=dw)))
A = torch.randn(8192, 4096, device=dev).bfloat16(); Bm = torch.randn(4096, 4096, device=dev).bfloat16()
o = torch.empty(8192, 4096, device=dev, dtype=torch.bfloat16)
t = timeit(lambda: ops.gemm_bf16(A, Bm, out=o, block_n=128))
print("gemm 8192x4096x4096 us", t, "TFLOP/s", 2 * 8192 * 4096 * 4096 / t / 1e6)
ref2 = (A[:256].float() @ Bm.float().t())
for bn in (128, 256):
    t = timeit(lambda: ops.gemm_bf16(A, Bm, out=o, block_n=bn))
    print("block_n", bn, "us", t, "TFLOP/s", 2 * 8192 * 4096 * 4096 / t / 1e6, "err", float((o[:256].float() - ref2).abs().max() / ref2.abs().max()))
A2 = torch.randn(16384, 8192, device=dev).bfloat16(); B2 = torch.randn(8192, 8192, device=dev).bfloat16()
o2 = torch.empty(16384, 8192, device=dev, dtype=torch.bfloat16)
for bn in (128, 256):
    t = timeit(lambda: ops.gemm_bf16(A2, B2, out=o2, block_n=bn), 5)
    print("16384x8192x8192 block_n", bn, "us", t, "TFLOP/s", 2 * 16384 * 8192 * 8192 / t / 1e6)
t = timeit(lambda: torch.matmul(A2, B2.t(), out=o2), 5)
print("cuBLAS 16384x8192x8192 us", t, "TFLOP/s", 2 * 16384 * 8192 * 8192 / t / 1e6)
for bn in (128, 256):
    t = timeit(lambda: ops.gemm_bf16(dy, x, a_mn=True, b_mn=True, out_dtype=torch.float32, out=dw, block_n=bn))
    print("fc1 wgrad block_n", bn, "us", t)
ref = dy.float().t() @ x.float()
print("ok", float((dw - ref).abs().max()))
