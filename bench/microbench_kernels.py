"""Per-kernel device time of the native ResNet-34 step shapes (batch 128), measured with CUDA events
around a CUDA graph of `reps` back-to-back launches (so launch overhead is excluded the same way the
training step excludes it).  Output: one line per op, microseconds per launch."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import ops
from kubeml_b200.ops import nn as K
from kubeml_b200.models.native import auto_splits, pick_block_n
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
REPS = 40
results = {}

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
    results[name] = round(best, 2)
    print(f"{best:8.2f} us  {name}", flush=True)

Nb = 128
for (tag, H, C) in (("layer1", 8, 64), ("layer2", 4, 128), ("layer3", 2, 256), ("layer4", 1, 512)):
    x = torch.randn(Nb, H, H, C, device=dev).bfloat16()
    w = torch.randn(C, 3, 3, C, device=dev).bfloat16()
    y = torch.empty(Nb, H, H, C, device=dev, dtype=torch.bfloat16)
    dw = torch.zeros(C, 3, 3, C, device=dev)
    M = Nb * H * H
    bn_ = pick_block_n((M + 127) // 128, C)
    timeit(f"{tag} fprop  M{M} C{C} bn{bn_}", lambda: ops.conv2d_fprop(x, w, 1, out=y, block_n=bn_))
    timeit(f"{tag} dgrad  M{M} C{C} bn{bn_}", lambda: ops.conv2d_dgrad(x, w, 1, H, H, out=y, block_n=bn_))
    tiles = ((M + 127) // 128) * ((C + bn_ - 1) // bn_)
    live = 9 if H > 1 else 1
    S = auto_splits(tiles, live * (C // 64), min_chunks=3)
    if S > 1:
        ys = torch.empty(S, Nb, H, H, C, device=dev, dtype=torch.float32)
        timeit(f"{tag} fprop  slabs S{S}", lambda: ops.conv2d_fprop(x, w, 1, out=ys, block_n=bn_, splits=S, slabs=True))
        timeit(f"{tag} dgrad  slabs S{S}", lambda: ops.conv2d_dgrad(x, w, 1, H, H, out=ys, block_n=bn_, splits=S, slabs=True))
    bw = 128 if C > 64 else 64
    tiles = ((C + 127) // 128) * ((C + bw - 1) // bw) * 9
    sp = auto_splits(tiles, (M + 63) // 64)
    timeit(f"{tag} wgrad  M{M} C{C} splits{sp}", lambda: ops.conv2d_wgrad(x, x, 3, 3, 1, out=dw, splits=sp, block_n=bw))
    xm = x.view(M, C); ym = y.view(M, C)
    gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    nbt = torch.zeros((), dtype=torch.long, device=dev)
    mean, istd = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    dg, db = torch.empty(C, device=dev), torch.empty(C, device=dev)
    dres = torch.empty_like(xm)
    timeit(f"{tag} bn_fwd M{M} C{C}", lambda: K.bn_fwd_train(xm, xm, ym, gamma, beta, rm, rv, nbt, mean, istd, True))
    timeit(f"{tag} bn_bwd M{M} C{C}", lambda: K.bn_bwd(xm, ym, xm, gamma, mean, istd, ym, dres, dg, db))
    if S > 1:
        xo = torch.empty_like(xm)
        ysm = ys.view(S, M, C)
        timeit(f"{tag} bn_fwd slabs S{S}", lambda: K.bn_fwd_train(None, xm, ym, gamma, beta, rm, rv, nbt, mean, istd, True, x_slabs=ysm, x_out=xo))
        timeit(f"{tag} bn_bwd slabs S{S}", lambda: K.bn_bwd(None, ym, xm, gamma, mean, istd, ym, dres, dg, db, dy_slabs=ysm))
# stem
xs = torch.randn(Nb, 32, 32, 3, device=dev).bfloat16()
cols = torch.empty(Nb * 256, 192, device=dev, dtype=torch.bfloat16)
wst = torch.randn(64, 192, device=dev).bfloat16()
c1 = torch.empty(Nb * 256, 64, device=dev, dtype=torch.bfloat16)
timeit("stem im2col", lambda: K.im2col(xs, cols, 7, 7, 2, 3))
timeit("stem gemm M32768 N64 K192", lambda: ops.gemm_bf16(cols, wst, out=c1))
dws = torch.zeros(64, 147, device=dev)
sp = auto_splits(2, 512)
timeit(f"stem wgrad splits{sp}", lambda: ops.gemm_bf16(c1, cols[:, :147], a_mn=True, b_mn=True, out=dws, splits=sp, block_n=128))
a1 = torch.randn(Nb, 16, 16, 64, device=dev).bfloat16(); p1 = torch.empty(Nb, 8, 8, 64, device=dev, dtype=torch.bfloat16)
idx = torch.empty(Nb, 8, 8, 64, device=dev, dtype=torch.uint8)
timeit("maxpool fwd", lambda: K.maxpool_fwd(a1, p1, idx, 3, 2, 1))
timeit("maxpool bwd", lambda: K.maxpool_bwd(p1, idx, a1, 3, 2, 1))
wflat = torch.randn(21_800_000 // 4 * 4, device=dev); gflat = torch.randn_like(wflat); wb = torch.empty_like(wflat, dtype=torch.bfloat16)
timeit("sgd 21.8M + bf16 copy", lambda: ops.sgd_step_(wflat, gflat, None, 0.1, 1e-4, w_bf16=wb))
lg = torch.randn(128, 1000, device=dev).bfloat16(); lb = torch.randint(0, 1000, (128,), device=dev)
timeit("softmax_ce 128x1000", lambda: ops.softmax_ce_raw(lg, lb))
timeit("fill 1 float4 (launch floor)", lambda: ops.fill_(dg[:4] if dg.numel() >= 4 else dg, 0.0))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(results, open("gpurun_out/microbench_kernels.json", "w"), indent=1)
