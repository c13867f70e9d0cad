"""Which native ops survive CUDA-graph capture?  (bring-up diagnostic)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import ops
from kubeml_b200.ops import nn as K
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)

def try_capture(name, fn, mode="global"):
    fn(); fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g, capture_error_mode=mode):
            fn()
        g.replay(); torch.cuda.synchronize()
        print(f"ok   {name} [{mode}]", flush=True)
    except Exception as e:
        print(f"FAIL {name} [{mode}]: {str(e).splitlines()[0][:160]}", flush=True)
        torch.cuda.synchronize()

M, C = 4096, 64
x = torch.randn(M, C, device=dev).bfloat16(); y = torch.empty_like(x)
gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
nbt = torch.zeros((), dtype=torch.long, device=dev)
mean, istd = torch.empty(C, device=dev), torch.empty(C, device=dev)
w = torch.randn(1 << 16, device=dev); g = torch.randn(1 << 16, device=dev)
A = torch.randn(256, 256, device=dev).bfloat16(); Bm = torch.randn(128, 256, device=dev).bfloat16()
out = torch.empty(256, 128, device=dev, dtype=torch.bfloat16)
xi = torch.randn(4, 8, 8, 64, device=dev).bfloat16(); wi = torch.randn(64, 3, 3, 64, device=dev).bfloat16()
yo = torch.empty(4, 8, 8, 64, device=dev, dtype=torch.bfloat16)
lg = torch.randn(128, 1000, device=dev).bfloat16(); lb = torch.randint(0, 1000, (128,), device=dev)

for mode in ("global", "thread_local"):
    try_capture("fill", lambda: ops.fill_(w, 0.0), mode)
    try_capture("sgd", lambda: ops.sgd_step_(w, g, None, 0.1, 1e-4), mode)
    try_capture("gemm", lambda: ops.gemm_bf16(A, Bm, out=out), mode)
    try_capture("conv", lambda: ops.conv2d_fprop(xi, wi, 1, out=yo), mode)
    try_capture("softmax_ce", lambda: ops.softmax_ce_raw(lg, lb), mode)
    try_capture("bn_coop", lambda: K.bn_fwd_train(x, None, y, gamma, beta, rm, rv, nbt, mean, istd, True), mode)
    try_capture("as_tensor", lambda: torch.as_tensor(type("S", (), {"__cuda_array_interface__": {"shape": (16,), "typestr": "<f4", "data": (w.data_ptr(), False), "version": 3, "strides": None}})(), device=dev).zero_(), mode)
