"""Short single-GPU program for `ncu --set full`: a few launches of each hot kernel at ResNet-34
step shapes (batch 128).  See profiles/README.md for the capture commands."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import _native, ops
from kubeml_b200.ops import nn as K
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
Nb = 128
for it in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    # layer2 conv 3x3 128->128 on 4x4 (M=2048): fprop as split-K slabs, dgrad, wgrad
    x = torch.randn(Nb, 4, 4, 128, device=dev).bfloat16()
    w = torch.randn(128, 3, 3, 128, device=dev).bfloat16()
    ys = torch.empty(4, Nb, 4, 4, 128, device=dev, dtype=torch.float32)
    ops.conv2d_fprop(x, w, 1, out=ys, block_n=64, splits=4, slabs=True)
    y = torch.empty(Nb, 4, 4, 128, device=dev, dtype=torch.bfloat16)
    ops.conv2d_dgrad(x, w, 1, 4, 4, out=y, block_n=64)
    dw = torch.zeros(128, 3, 3, 128, device=dev)
    ops.conv2d_wgrad(x, x, 3, 3, 1, out=dw, splits=16, block_n=128)
    # round 2: convolution + BatchNorm (+ReLU) in one launch, and data gradient + BatchNorm backward in one
    # launch (kml_bn_fuse), at the layer2 shape the models use; a stride-2 convolution without im2col
    from kubeml_b200.models.native import fused_plan, pick_block_n
    C = 128
    gam, bet = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    sm_, si_ = torch.empty(C, device=dev), torch.empty(C, device=dev)
    raw = torch.empty(Nb, 4, 4, C, device=dev, dtype=torch.bfloat16)
    yf = torch.empty_like(raw)
    bn_, S_ = fused_plan(dev, 16, C, 18, pick_block_n(16, C), False)
    ops.conv2d_fprop(x, w, 1, out=yf, block_n=bn_, splits=S_, cluster=True,
                     bn_fuse=ops.BnFwd(gam, bet, rm, rv, None, sm_, si_, raw, relu=True))
    dgm, dbt = torch.empty(C, device=dev), torch.empty(C, device=dev)
    dxf, gf_ = torch.empty_like(raw), torch.empty_like(raw)
    bn_, S_ = fused_plan(dev, 16, C, 18, pick_block_n(16, C), True)
    ops.conv2d_dgrad(x, w, 1, 4, 4, addend=raw, out=dxf, block_n=bn_, splits=S_, cluster=True,
                     bn_fuse=ops.BnBwd(gam, sm_, si_, raw, dgm, dbt, mask_y=yf, raw_out=gf_))
    x8 = torch.randn(Nb, 8, 8, 64, device=dev).bfloat16()
    w2 = torch.randn(128, 3, 3, 64, device=dev).bfloat16()
    ops.conv2d_fprop(x8, w2, 1, stride=2, splits=4, cluster=True, out=torch.empty(Nb, 4, 4, 128, device=dev, dtype=torch.bfloat16))
    ops.conv2d_wgrad(x8, raw, 3, 3, 1, stride=2, splits=8, out=torch.zeros(128, 3, 3, 64, device=dev))
    # a large plain GEMM so the tensor pipe is visible: 8192 x 4096 x 4096
    A = torch.randn(8192, 4096, device=dev).bfloat16(); Bm = torch.randn(4096, 4096, device=dev).bfloat16()
    o = torch.empty(8192, 4096, device=dev, dtype=torch.bfloat16)
    ops.gemm_bf16(A, Bm, out=o, block_n=128)
    # the same convolution as a cluster split-K GEMM with BatchNorm statistics, then the element-wise BatchNorm
    stc = torch.zeros(K.bn_stats_numel(128), device=dev)
    yc = torch.empty(Nb, 4, 4, 128, device=dev, dtype=torch.bfloat16)
    ops.conv2d_fprop(x, w, 1, out=yc, block_n=64, splits=8, cluster=True, stats=stc)
    g1, b1 = torch.ones(128, device=dev), torch.zeros(128, device=dev)
    mn, sd = torch.zeros(128, device=dev), torch.ones(128, device=dev)
    K.bn_apply_train(yc, stc, None, torch.empty_like(yc), g1, b1, None, None, None, mn, sd, True)
    # two-SM GEMM (cta_group::2) and the persistent one-CTA GEMM on the large problem
    A2 = torch.randn(8192, 4096, device=dev).bfloat16(); B2 = torch.randn(4096, 4096, device=dev).bfloat16()
    o2 = torch.empty(8192, 4096, device=dev, dtype=torch.bfloat16)
    ops.gemm_bf16(A2, B2, out=o2, block_n=512)
    # BN forward on slabs, backward
    M, C = 2048, 128
    gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    mean, istd = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    yb = torch.empty(M, C, device=dev, dtype=torch.bfloat16); xo = torch.empty_like(yb)
    K.bn_fwd_train(None, None, yb, gamma, beta, None, None, None, mean, istd, True, x_slabs=ys.view(4, M, C), x_out=xo)
    dx = torch.empty_like(yb); dg, db = torch.empty(C, device=dev), torch.empty(C, device=dev)
    K.bn_bwd(yb, yb, xo, gamma, mean, istd, dx, None, dg, db)
    # K-AVG merge, two replicas on one GPU (single-launch emulation), ResNet-34 size
    n = 21_797_888
    reps = [torch.randn(n, device=dev) for _ in range(2)]
    p = _native.MergeParams()
    for r in range(2):
        p.replica[r] = reps[r].data_ptr()
    p.n_f32, p.my_rank, p.participants, p.write_mask, p.generation = n, 0, 3, 3, 0
    st = torch.zeros(1, dtype=torch.int32, device=dev); p.status = st.data_ptr()
    _native.check(_native.lib().kml_kavg_merge(ctypes.byref(p), _native.stream_ptr()), "merge")
    # optimizer and loss
    wf = torch.randn(n, device=dev); gf = torch.randn(n, device=dev); wb = torch.empty(n, device=dev, dtype=torch.bfloat16)
    ops.sgd_step_(wf, gf, None, 0.1, 1e-4, w_bf16=wb)
    lg = torch.randn(128, 1000, device=dev).bfloat16(); lb = torch.randint(0, 1000, (128,), device=dev)
    ops.softmax_ce_raw(lg, lb)
    torch.cuda.synchronize()
print("ok")
