"""Fixed-cost floor of the kernel structures (bring-up diagnostic)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import ops
from kubeml_b200.ops import nn as K
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
REPS = 40
def timeit(name, fn):
    fn(); fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(REPS): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(5):
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / REPS)
    print(f"{best:8.2f} us  {name}", flush=True)
for (M, N, Kd) in ((128, 64, 64), (128, 64, 256), (128, 64, 512), (128, 64, 1024), (128, 128, 64), (1024, 64, 64), (128 * 148, 64, 64)):
    A = torch.randn(M, Kd, device=dev).bfloat16(); Bm = torch.randn(N, Kd, device=dev).bfloat16()
    o = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    timeit(f"gemm M{M} N{N} K{Kd}", lambda: ops.gemm_bf16(A, Bm, out=o))
for (M, C) in ((8, 64), (128, 64), (128, 512), (8192, 64)):
    x = torch.randn(M, C, device=dev).bfloat16(); y = torch.empty_like(x)
    gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    mean, istd = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    timeit(f"bn_fwd_train M{M} C{C} (cooperative)", lambda: K.bn_fwd_train(x, None, y, gamma, beta, None, None, None, mean, istd, True))
    timeit(f"bn_fwd_eval  M{M} C{C} (plain launch)", lambda: K.bn_fwd_eval(x, None, y, gamma, beta, mean, istd, True))
w = torch.randn(1 << 12, device=dev); g = torch.randn(1 << 12, device=dev)
timeit("sgd 4K (plain launch)", lambda: ops.sgd_step_(w, g, None, 0.1, 1e-4))
