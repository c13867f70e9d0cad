"""One small launch of every kernel family, for compute-sanitizer (scripts/sanitize.sh):
  compute-sanitizer --tool memcheck|racecheck|synccheck python bench/sanitize_target.py
Shapes are tiny on purpose (the tools slow kernels down 10-100x) but cover every code path: one-CTA GEMM
(K-/MN-major operands, split-K atomics, slabs), cluster split-K with statistics / mask / BatchNorm sums,
persistent many-tile launch, two-SM pair kernel, implicit-GEMM conv forward / dgrad / wgrad, BatchNorm
(grid-barrier and element-wise forms), pooling, im2col/col2im, VGG layer kernels, loss, optimizers, merge."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import _native, ops
from kubeml_b200.ops import nn as K
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
bf = lambda *s: torch.randn(*s, device=dev).bfloat16()
f32 = lambda *s: torch.randn(*s, device=dev)

# ---- GEMM family
a, b = bf(200, 320), bf(96, 320)
ops.gemm_bf16(a, b, bias=f32(96), relu=True)
ops.gemm_bf16(a, b, out_dtype=torch.float32, splits=3)
ops.gemm_bf16(a, b, out=torch.empty(3, 200, 96, device=dev), splits=3, slabs=True)
ops.gemm_bf16(bf(320, 200), bf(320, 96), a_mn=True, b_mn=True, out_dtype=torch.float32)
ops.gemm_bf16(a, bf(320, 96), b_mn=True)
st = torch.zeros(K.bn_stats_numel(96), device=dev)
ops.gemm_bf16(a, b, out=torch.empty(200, 96, device=dev, dtype=torch.bfloat16), splits=4, cluster=True, stats=st)
ops.gemm_bf16(bf(128 * 700, 64), bf(64, 64))                         # > 592 tiles: persistent
ops.gemm_bf16(bf(640, 200), bf(304, 200), bias=f32(304), relu=True, block_n=512)   # two-SM pair kernel
ops.gemm_bf16(bf(256, 128), bf(512, 128), block_n=256)

# ---- convolution family
x, w = bf(8, 8, 8, 64), bf(64, 3, 3, 64) * 0.1
y = ops.conv2d_fprop(x, w, 1, bias=f32(64), relu=True)
st2 = torch.zeros(K.bn_stats_numel(64), device=dev)
yc = torch.empty(8, 8, 8, 64, device=dev, dtype=torch.bfloat16)
ops.conv2d_fprop(x, w, 1, out=yc, block_n=64, splits=3, cluster=True, stats=st2)
mean, istd = f32(64), torch.rand(64, device=dev) + 0.5
g = torch.empty_like(x)
bst = torch.zeros(K.bn_stats_numel(64), device=dev)
ops.conv2d_dgrad(bf(8, 8, 8, 64), w, 1, 8, 8, addend=bf(8, 8, 8, 64), out=g, block_n=64, splits=3, cluster=True,
                 mask_y=x, bn=(x, mean, istd), stats=bst)
ops.conv2d_dgrad(bf(8, 8, 8, 64), w, 1, 8, 8)
ops.conv2d_wgrad(x, bf(8, 8, 8, 64), 3, 3, 1, splits=2)

# ---- BatchNorm
gamma, beta = torch.ones(64, device=dev), torch.zeros(64, device=dev)
rm, rv = torch.zeros(64, device=dev), torch.ones(64, device=dev)
nbt = torch.zeros((), dtype=torch.long, device=dev)
yb = torch.empty_like(x)
K.bn_apply_train(yc, st2, x, yb, gamma, beta, rm, rv, nbt, mean, istd, True)
K.bn_fwd_train(x, None, yb, gamma, beta, rm, rv, nbt, mean, istd, True)
sl = torch.randn(3, 8, 8, 8, 64, device=dev)
xo = torch.empty_like(x)
K.bn_fwd_train(None, x, yb, gamma, beta, rm, rv, nbt, mean, istd, True, x_slabs=sl, x_out=xo)
K.bn_fwd_eval(x, None, yb, gamma, beta, rm, rv, True)
dx, dres = torch.empty_like(x), torch.empty_like(x)
dg, db = torch.empty(64, device=dev), torch.empty(64, device=dev)
K.bn_bwd(bf(8, 8, 8, 64), yb, x, gamma, mean, istd, dx, dres, dg, db)
K.bn_bwd_apply(g, x, gamma, mean, istd, bst, dx, dg, db)

# ---- pooling, im2col, VGG layer kernels
p = torch.empty(8, 4, 4, 64, device=dev, dtype=torch.bfloat16); idx = torch.empty(8, 4, 4, 64, device=dev, dtype=torch.uint8)
K.maxpool_fwd(x, p, idx, 3, 2, 1); K.maxpool_bwd(p, idx, dx, 3, 2, 1)
av = torch.empty(8, 64, device=dev, dtype=torch.bfloat16); K.avgpool_fwd(x, av); K.avgpool_bwd(av, dx)
cols = torch.empty(8 * 64, 576, device=dev, dtype=torch.bfloat16); K.im2col(x, cols, 3, 3, 1, 1); K.col2im(cols, x, dx, 3, 3, 1, 1)
x3 = bf(4, 9, 9, 3); c3 = torch.empty(4 * 81, 64, device=dev, dtype=torch.bfloat16); K.im2col(x3, c3, 3, 3, 1, 1)
dbias = torch.empty(64, device=dev)
K.relu_bwd_cols(bf(512, 64), x.view(512, 64), g.view(512, 64), 2.0, dbias)
K.relu_bwd_cols(None, x.view(512, 64), g.view(512, 64), 1.0, dbias, dy_slabs=sl.view(3, 512, 64))
K.bias_act(sl.view(3, 512, 64), dbias, g.view(512, 64), True)
K.colsum(bf(100, 24), torch.empty(24, device=dev))
cnt = torch.zeros(1, dtype=torch.int64, device=dev)
K.dropout_fwd(x, g, None, cnt, 1, 0.5); K.bump_counter(cnt)
rep = torch.empty(8, 64 * 49, device=dev, dtype=torch.bfloat16); K.replicate_fwd(av, rep, 49); K.replicate_bwd(rep, av, 49)
img = torch.randint(0, 255, (8, 32, 32, 3), device=dev, dtype=torch.uint8)
K.image_prep(img, torch.empty(8, 32, 32, 3, device=dev, dtype=torch.bfloat16), None, 4, (0.5, 0.5, 0.5), (0.2, 0.2, 0.2))

# ---- loss, optimizers, merge
lg = bf(64, 100); lb = torch.randint(0, 100, (64,), device=dev)
ops.softmax_ce_raw(lg, lb, want_grad=True, want_correct=True)
n = 1 << 14
wv, gv, m1, m2 = f32(n), f32(n), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
wb = torch.empty(n, device=dev, dtype=torch.bfloat16)
ops.sgd_step_(wv, gv, m1, 0.1, 1e-4, 0.9, False, w_bf16=wb)
ops.adam_step_(wv, gv, m1, m2, 1, 1e-3, w_bf16=wb)
ops.adam_step_(wv, gv, m1, m2, 0, 1e-3, step_dev=torch.zeros(2, dtype=torch.int32, device=dev))
reps = [f32(n) for _ in range(2)]
mp = _native.MergeParams()
for r in range(2):
    mp.replica[r] = reps[r].data_ptr()
mp.n_f32, mp.my_rank, mp.participants, mp.write_mask, mp.generation = n, 0, 3, 3, 0
status = torch.zeros(1, dtype=torch.int32, device=dev); mp.status = status.data_ptr()
_native.check(_native.lib().kml_kavg_merge(ctypes.byref(mp), _native.stream_ptr()), "merge")
torch.cuda.synchronize()
print("sanitize target ok")
