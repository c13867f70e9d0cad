"""GPU numerics: every hand-written kernel against a plain PyTorch fp32 restatement of the same op."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    d = torch.device("cuda", 0)
    torch.cuda.set_device(d)
    from kubeml_b200 import _native
    _native.lib()           # the extension must be the thing under test: fail loudly if absent
    return d


def test_native_library_loaded(dev):
    from kubeml_b200 import _native
    assert _native.lib().kml_version() >= 100


# ------------------------------------------------------------------------------------- optimizers
@pytest.mark.parametrize("momentum,nesterov,wd", [(0.0, False, 1e-4), (0.9, False, 1e-4), (0.9, True, 0.0)])
def test_sgd_matches_torch(dev, momentum, nesterov, wd):
    from kubeml_b200 import ops
    torch.manual_seed(0)
    n = 1 << 16
    w0 = torch.randn(n, device=dev)
    p = torch.nn.Parameter(w0.clone())
    opt = torch.optim.SGD([p], lr=0.05, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    w = w0.clone()
    mb = torch.zeros(n, device=dev)
    wb = torch.empty(n, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        g = torch.randn(n, device=dev)
        p.grad = g.clone()
        opt.step()
        ops.sgd_step_(w, g, mb if momentum else None, 0.05, wd, momentum, nesterov, w_bf16=wb)
    torch.testing.assert_close(w, p.data, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(wb.float(), w.bfloat16().float(), rtol=0, atol=0)


def test_adam_matches_torch(dev):
    from kubeml_b200 import ops
    torch.manual_seed(0)
    n = 1 << 14
    w0 = torch.randn(n, device=dev)
    p = torch.nn.Parameter(w0.clone())
    opt = torch.optim.Adam([p], lr=1e-2, weight_decay=1e-4)
    w, m, v = w0.clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    for t in range(1, 4):
        g = torch.randn(n, device=dev)
        p.grad = g.clone()
        opt.step()
        ops.adam_step_(w, g, m, v, t, 1e-2, (0.9, 0.999), 1e-8, 1e-4)
    torch.testing.assert_close(w, p.data, rtol=2e-5, atol=2e-6)


# ------------------------------------------------------------------------------------- softmax-CE
@pytest.mark.parametrize("rows,classes,dtype", [(128, 1000, torch.float32), (64, 10, torch.float32),
                                                (37, 100, torch.bfloat16), (1, 10, torch.float32)])
def test_softmax_ce(dev, rows, classes, dtype):
    from kubeml_b200 import ops
    torch.manual_seed(1)
    x = (torch.randn(rows, classes, device=dev) * 3).to(dtype)
    y = torch.randint(0, classes, (rows,), device=dev)
    xr = x.float().clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(xr, y)
    ref.backward()
    for _ in range(2):      # twice: the kernel must leave its scratch reusable
        loss, dl, corr = ops.softmax_ce_raw(x, y, want_grad=True, want_correct=True)
        tol = 1e-5 if dtype == torch.float32 else 2e-2
        torch.testing.assert_close(loss.view(()), ref.detach(), rtol=tol, atol=tol)
        torch.testing.assert_close(dl.float(), xr.grad, rtol=tol, atol=tol if dtype == torch.float32 else 1e-3)
        assert int(corr.item()) == int(x.float().argmax(1).eq(y).sum().item())


def test_cross_entropy_autograd(dev):
    from kubeml_b200 import ops
    torch.manual_seed(2)
    x = torch.randn(32, 10, device=dev, requires_grad=True)
    y = torch.randint(0, 10, (32,), device=dev)
    ops.cross_entropy(x, y).backward()
    g = x.grad.clone()
    x.grad = None
    torch.nn.functional.cross_entropy(x, y).backward()
    torch.testing.assert_close(g, x.grad, rtol=1e-5, atol=1e-6)


# ------------------------------------------------------------------------------------- K-AVG merge
def _merge_single_gpu(dev, replicas_f, replicas_i, parts, write, fuse=None):
    """All 'ranks' live on one GPU; one launch with generation 0 reduces the whole buffer."""
    from kubeml_b200 import _native
    lib = _native.lib()
    p = _native.MergeParams()
    for r, t in enumerate(replicas_f):
        p.replica[r] = t.data_ptr()
        if replicas_i is not None:
            p.ireplica[r] = replicas_i[r].data_ptr()
    p.n_f32 = replicas_f[0].numel()
    p.n_i64 = replicas_i[0].numel() if replicas_i is not None else 0
    p.my_rank = parts[0]
    p.participants = sum(1 << r for r in parts)
    p.write_mask = sum(1 << r for r in write)
    p.generation = 0
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    p.status = status.data_ptr()
    if fuse is not None:
        p.fuse_sgd = 1
        p.grad = fuse["grad"].data_ptr()
        p.n_sgd = fuse["n"]
        p.n_fuse = fuse.get("n_fuse", 0)
        p.lr, p.weight_decay, p.momentum = fuse["lr"], fuse["wd"], fuse["mom"]
        p.momentum_buf = fuse["mbuf"].data_ptr() if fuse.get("mbuf") is not None else None
        if fuse.get("wbf16") is not None:
            p.w_bf16 = fuse["wbf16"].data_ptr()
        if fuse.get("adam") is not None:          # (exp_avg_sq, beta1, beta2, eps, step_dev)
            p.fuse_sgd = 2
            p.state2 = fuse["adam"][0].data_ptr()
            p.beta1, p.beta2, p.eps = fuse["adam"][1], fuse["adam"][2], fuse["adam"][3]
            p.step_dev = fuse["adam"][4].data_ptr()
    _native.check(lib.kml_kavg_merge(ctypes.byref(p), _native.stream_ptr()), "kml_kavg_merge")
    torch.cuda.synchronize()
    assert int(status.item()) == 0


@pytest.mark.parametrize("nranks,parts", [(2, [0, 1]), (4, [0, 1, 2, 3]), (8, list(range(8))), (4, [0, 2]), (3, [1])])
def test_merge_matches_cpu_oracle(dev, nranks, parts):
    from kubeml_b200.model.model import kavg_oracle
    torch.manual_seed(3)
    n = 4 * 12345
    reps = [torch.randn(n, device=dev) for _ in range(nranks)]
    ints = [torch.randint(0, 1000, (7,), device=dev, dtype=torch.int64) for _ in range(nranks)]
    want = kavg_oracle([{"w": reps[r].cpu().numpy(), "c": ints[r].cpu().numpy()} for r in parts])
    untouched = {r: reps[r].clone() for r in range(nranks)}
    write = list(range(nranks)) if len(parts) > 1 else parts
    _merge_single_gpu(dev, reps, ints, parts, write)
    for r in write:
        # same summation order (ascending rank) and a true division: bit-exact
        np.testing.assert_array_equal(reps[r].cpu().numpy(), want["w"])
        np.testing.assert_array_equal(ints[r].cpu().numpy(), want["c"])
    for r in set(range(nranks)) - set(write):
        torch.testing.assert_close(reps[r], untouched[r], rtol=0, atol=0)


def test_merge_fused_last_sgd_step_and_reset(dev):
    torch.manual_seed(4)
    n, npar = 4 * 4096, 4 * 3000
    w = [torch.randn(n, device=dev) for _ in range(2)]
    g = torch.randn(npar, device=dev)
    mb = torch.randn(npar, device=dev)
    w0 = w[0].clone()
    lr, wd, mom = 0.1, 1e-4, 0.9
    d = g + wd * w0[:npar]
    buf = mom * mb + d
    stepped = w0.clone()
    stepped[:npar] -= lr * buf
    want = (stepped + w[1]) / 2
    _merge_single_gpu(dev, w, None, [0, 1], [0, 1],
                      fuse=dict(grad=g, n=npar, lr=lr, wd=wd, mom=mom, mbuf=mb))
    torch.testing.assert_close(w[0], want, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(w[1], want, rtol=1e-6, atol=1e-6)
    assert float(mb.abs().max()) == 0.0       # optimizer state reset in the same pass


def test_merge_single_participant_is_the_fused_step_only(dev):
    """One participant, one destination: the average is the identity, so the kernel is exactly the round's
    last SGD step (+ bf16 copy + optimizer reset) - what a 1-GPU job pays per sync."""
    torch.manual_seed(6)
    n, npar = 4 * 5000, 4 * 4000
    w = torch.randn(n, device=dev)
    g = torch.randn(npar, device=dev)
    mb = torch.randn(npar, device=dev)
    wb = torch.zeros(npar, dtype=torch.bfloat16, device=dev)
    lr, wd, mom = 0.05, 1e-3, 0.9
    want = w.clone()
    want[:npar] -= lr * (mom * mb + g + wd * w[:npar])
    _merge_single_gpu(dev, [w], None, [0], [0], fuse=dict(grad=g, n=npar, lr=lr, wd=wd, mom=mom, mbuf=mb, wbf16=wb))
    torch.testing.assert_close(w, want, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(wb.float(), want[:npar].bfloat16().float(), rtol=0, atol=0)
    assert float(mb.abs().max()) == 0.0
    # without a pending step nothing changes
    before = w.clone()
    _merge_single_gpu(dev, [w], None, [0], [0])
    torch.testing.assert_close(w, before, rtol=0, atol=0)


@pytest.mark.parametrize("nrep", [1, 2])
def test_merge_fuses_only_the_head_when_the_tail_was_stepped_early(dev, nrep):
    """`n_fuse`: the caller already applied the step to [n_fuse, n_sgd) under the backward pass (layer4 + fc of the
    ResNets); the merge kernel's phase 0 touches [0, n_fuse) only - weights, bf16 copy - and still resets the whole
    optimizer state."""
    torch.manual_seed(16)
    n, npar, head = 4 * 5000, 4 * 4000, 4 * 1500
    w = [torch.randn(n, device=dev) for _ in range(nrep)]
    g = torch.randn(npar, device=dev)
    mb = torch.randn(npar, device=dev)
    wb = torch.full((npar,), 7.0, dtype=torch.bfloat16, device=dev)
    lr, wd, mom = 0.05, 1e-3, 0.9
    stepped = w[0].clone()
    stepped[:head] -= lr * (mom * mb[:head] + g[:head] + wd * w[0][:head])
    want = stepped.clone()
    for other in w[1:]:
        want = want + other
    want = want / nrep
    parts = list(range(nrep))
    _merge_single_gpu(dev, w, None, parts, parts,
                      fuse=dict(grad=g, n=npar, n_fuse=head, lr=lr, wd=wd, mom=mom, mbuf=mb, wbf16=wb))
    for r in parts:
        torch.testing.assert_close(w[r], want, rtol=1e-6, atol=1e-6)
    assert float(mb.abs().max()) == 0.0
    if nrep == 1:       # alone, the kernel writes the bf16 copy of what it stepped and leaves the rest to the caller
        torch.testing.assert_close(wb[:head].float(), want[:head].bfloat16().float(), rtol=0, atol=0)
        assert bool((wb[head:].float() == 7.0).all())
    else:               # after a real merge the whole copy is refreshed
        torch.testing.assert_close(wb.float(), want[:npar].bfloat16().float(), rtol=0, atol=0)


@pytest.mark.parametrize("nrep", [1, 2])
def test_merge_fused_last_adam_step_and_reset(dev, nrep):
    """The Adam form of the fused last step (function_vgg11.py's optimizer): replica 0 takes its third Adam step
    inside the merge kernel, then the replicas are averaged and both moments and the step count are reset."""
    torch.manual_seed(8)
    n = 4 * 2048
    p0 = torch.randn(n, device=dev, requires_grad=True)
    opt = torch.optim.Adam([p0], lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=1e-3)
    grads = [torch.randn(n, device=dev) for _ in range(3)]
    for g in grads[:2]:                      # two ordinary steps build up the moments
        p0.grad = g.clone(); opt.step()
    st = opt.state[p0]
    w = [p0.detach().clone()] + [torch.randn(n, device=dev) for _ in range(nrep - 1)]
    m1, m2 = st["exp_avg"].clone(), st["exp_avg_sq"].clone()
    step_dev = torch.tensor([2, 0], dtype=torch.int32, device=dev)
    p0.grad = grads[2].clone(); opt.step()   # the oracle's third step
    want = p0.detach().clone()
    for other in w[1:]:
        want = want + other
    want = want / nrep
    parts = list(range(nrep))
    _merge_single_gpu(dev, w, None, parts, parts,
                      fuse=dict(grad=grads[2], n=n, lr=1e-2, wd=1e-3, mom=0.0, mbuf=m1, adam=(m2, 0.9, 0.99, 1e-8, step_dev)))
    for r in parts:
        torch.testing.assert_close(w[r], want, rtol=2e-5, atol=2e-6)
    assert float(m1.abs().max()) == 0.0 and float(m2.abs().max()) == 0.0
    assert int(step_dev[0].item()) == 0


# ------------------------------------------------------------------------------------- tcgen05 GEMM
def _rel_err(got, want):
    return float((got.float() - want).abs().max() / want.abs().max().clamp_min(1e-6))


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (128, 128, 256), (256, 512, 512), (100, 72, 200),
                                   (1024, 1000, 512), (128, 10, 512)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True)])
def test_gemm_bf16(dev, M, N, K, a_mn, b_mn):
    from kubeml_b200 import ops
    if (a_mn and M % 8) or (b_mn and N % 8) or K % 8 or ((not b_mn) and K % 8):
        pytest.skip("leading dimension not a multiple of 8")
    torch.manual_seed(5)
    A = torch.randn(M, K, device=dev).bfloat16()
    B = torch.randn(N, K, device=dev).bfloat16()
    want = A.float() @ B.float().t()
    a = A.t().contiguous() if a_mn else A
    b = B.t().contiguous() if b_mn else B
    got = ops.gemm_bf16(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32)
    torch.cuda.synchronize()
    assert _rel_err(got, want) < 2e-3


def test_gemm_bias_relu_bf16_out_and_splitk(dev):
    from kubeml_b200 import ops
    torch.manual_seed(6)
    M, N, K = 256, 192, 1024
    A = torch.randn(M, K, device=dev).bfloat16()
    B = torch.randn(N, K, device=dev).bfloat16()
    bias = torch.randn(N, device=dev)
    want = torch.relu(A.float() @ B.float().t() + bias)
    got = ops.gemm_bf16(A, B, bias=bias, relu=True)
    assert got.dtype == torch.bfloat16 and _rel_err(got, want) < 1e-2
    got2 = ops.gemm_bf16(A, B, out_dtype=torch.float32, splits=4)
    assert _rel_err(got2, A.float() @ B.float().t()) < 2e-3


@pytest.mark.parametrize("N,H,C,Co,k,pad", [(4, 8, 64, 64, 3, 1), (16, 4, 128, 128, 3, 1), (64, 2, 256, 256, 3, 1),
                                            (128, 1, 512, 512, 3, 1), (2, 16, 64, 128, 3, 1), (8, 8, 64, 64, 1, 0),
                                            (3, 8, 128, 72, 3, 1)])
def test_conv2d_fprop(dev, N, H, C, Co, k, pad):
    from kubeml_b200 import ops
    torch.manual_seed(7)
    x = torch.randn(N, H, H, C, device=dev).bfloat16()
    w = (torch.randn(Co, k, k, C, device=dev) / (k * k * C) ** 0.5).bfloat16()
    want = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2),
                                      padding=pad).permute(0, 2, 3, 1)
    got = ops.conv2d_fprop(x, w, pad, out_dtype=torch.float32)
    torch.cuda.synchronize()
    assert got.shape == want.shape
    assert _rel_err(got, want) < 2e-3


@pytest.mark.parametrize("N,H,C,Co,k,pad,splits", [(4, 8, 64, 64, 3, 1, 1), (16, 4, 128, 128, 3, 1, 2),
                                                   (128, 1, 512, 512, 3, 1, 1), (32, 8, 64, 64, 3, 1, 8),
                                                   (8, 8, 64, 128, 1, 0, 1), (6, 2, 256, 72, 3, 1, 1)])
def test_conv2d_wgrad(dev, N, H, C, Co, k, pad, splits):
    from kubeml_b200 import ops
    torch.manual_seed(8)
    x = torch.randn(N, H, H, C, device=dev).bfloat16()
    Ho = H + 2 * pad - k + 1
    dy = torch.randn(N, Ho, Ho, Co, device=dev).bfloat16()
    want = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (Co, C, k, k),
                                       dy.float().permute(0, 3, 1, 2), padding=pad).permute(0, 2, 3, 1)
    got = ops.conv2d_wgrad(x, dy, k, k, pad, splits=splits)
    torch.cuda.synchronize()
    assert _rel_err(got, want) < 2e-3


# ------------------------------------------------------------------------------------- arena
def test_arena_bind_and_flat_optimizer(dev):
    from kubeml_b200.optim import SGD
    from kubeml_b200.parallel.arena import ArenaLayout, WeightArena

    def mk():
        torch.manual_seed(1)
        return torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3), torch.nn.BatchNorm2d(8), torch.nn.Flatten(),
                                   torch.nn.Linear(8 * 36, 10)).to(dev)
    m1, m2 = mk(), mk()
    arena = WeightArena(ArenaLayout(m2), dev)
    arena.bind(m2)
    o1 = torch.optim.SGD(m1.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    o2 = SGD(m2.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    o2.attach_arena(arena)
    x = torch.randn(4, 3, 8, 8, device=dev)
    y = torch.randint(0, 10, (4,), device=dev)
    for _ in range(3):
        for m, o in ((m1, o1), (m2, o2)):
            o.zero_grad()
            torch.nn.functional.cross_entropy(m(x), y).backward()
            o.step()
    assert o2._flat_views() is not None
    for a, b in zip(m1.state_dict().values(), m2.state_dict().values()):
        torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-5)
    assert int(arena.counters[0].item()) == 3
    arena.close()


# ------------------------------------------------------------------------------------- two-SM GEMM
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (1024, 768, 4096), (640, 304, 200), (128, 256, 128)])
def test_gemm_two_sm_cta_pair(dev, M, N, K):
    """tcgen05.mma.cta_group::2 kernel (block_n=512 selects it): bias + ReLU bf16 output and plain fp32
    output against fp32 torch, including ragged M / N / K and an odd number of 128-row tiles."""
    from kubeml_b200 import ops
    torch.manual_seed(0)
    a = torch.randn(M, K, device=dev).bfloat16()
    b = (torch.randn(N, K, device=dev) * 0.1).bfloat16()
    bias = torch.randn(N, device=dev)
    ref = a.float() @ b.float().t()
    out = torch.full((M, N), 7.0, device=dev, dtype=torch.bfloat16)
    ops.gemm_bf16(a, b, bias=bias, relu=True, out=out, block_n=512)
    want = torch.relu(ref + bias)
    assert float((out.float() - want).abs().max()) <= 1e-2 * float(want.abs().max())
    of = torch.empty(M, N, device=dev)
    ops.gemm_bf16(a, b, out=of, block_n=512)
    torch.cuda.synchronize()
    assert float((of - ref).abs().max()) <= 2e-3 * float(ref.abs().max())


def test_every_kernel_family_runs(dev):
    """bench/sanitize_target.py launches every kernel family once on small shapes (the compute-sanitizer
    target of scripts/sanitize.sh); it must at least run clean and leave the device healthy."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench", "sanitize_target.py")], capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0 and "sanitize target ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    assert torch.zeros(4, device=dev).sum().item() == 0.0
