"""The native model *programs* (forward + hand-ordered backward) against torchvision, run through
the CPU fp32 restatement of every kernel: exact agreement here + per-kernel GPU numerics
(tests/test_gpu_layers.py) is what pins the GPU path."""
import numpy as np
import pytest
import torch

from kubeml_b200 import ops
from kubeml_b200.models import native as nv


@pytest.fixture()
def fp32_activations():
    nv.set_activation_dtype(torch.float32)
    yield
    nv.set_activation_dtype(torch.bfloat16)


@pytest.mark.parametrize("bn_bwd_fusion", ["0", "1"])
def test_resnet18_program_matches_torchvision(fp32_activations, monkeypatch, bn_bwd_fusion):
    import torchvision
    monkeypatch.setenv("KUBEML_CLUSTER_BNBWD", bn_bwd_fusion)
    from kubeml_b200.models import resnet18
    torch.manual_seed(0)
    net = resnet18(num_classes=10)
    net._ensure_arena()
    with torch.no_grad():
        for p in net.parameters():
            if p.dim() >= 2:
                p.copy_(p.bfloat16().float())        # the tensor-core copy is bf16: make it lossless
    net._arena.refresh_bf16()
    ref = torchvision.models.resnet18(num_classes=10)
    assert list(net.state_dict().keys()) == list(ref.state_dict().keys())
    assert all(a.shape == b.shape for a, b in zip(net.state_dict().values(), ref.state_dict().values()))
    ref.load_state_dict(net.state_dict())
    net.train(); ref.train()
    x = torch.randn(8, 3, 32, 32)
    y = torch.randint(0, 10, (8,))
    out = net(x)
    ops.cross_entropy(out, y).backward()
    out_r = ref(x)
    torch.nn.functional.cross_entropy(out_r, y).backward()
    torch.testing.assert_close(out, out_r, rtol=1e-4, atol=1e-4)
    for n, p in ref.named_parameters():
        g = net._arena.grad(n)
        scale = p.grad.abs().max().item() + 1e-12
        assert (g - p.grad).abs().max().item() / scale < 1e-3, n
    torch.testing.assert_close(net.bn1.running_mean, ref.bn1.running_mean, rtol=1e-5, atol=1e-6)
    assert int(net.layer4[1].bn2.num_batches_tracked) == 1
    # a second step accumulates nothing stale (split-K gradients are cleared by the backward itself)
    net(x).float().sum().backward()
    ref(x).sum().backward()
    net.eval(); ref.eval()
    with torch.no_grad():
        torch.testing.assert_close(net(x), ref(x), rtol=1e-3, atol=1e-3)


def test_filter_is_stored_ohwi_but_exposed_oihw():
    from kubeml_b200.models import resnet18
    net = resnet18(num_classes=10)
    net._ensure_arena()
    w = net.layer1[0].conv1.weight
    assert tuple(w.shape) == (64, 64, 3, 3) and not w.is_contiguous()
    phys = net._arena.physical("layer1.0.conv1.weight")
    assert tuple(phys.shape) == (64, 3, 3, 64)
    with torch.no_grad():
        w[5, 7, 1, 2] = 42.0
    assert float(phys[5, 1, 2, 7]) == 42.0
    sd = net.state_dict()
    assert tuple(sd["layer1.0.conv1.weight"].shape) == (64, 64, 3, 3)
    net2 = resnet18(num_classes=10)
    net2.load_state_dict(sd)
    torch.testing.assert_close(net2.layer1[0].conv1.weight, w)


def test_arena_layout_flat_optimizer_and_bf16_copy():
    from kubeml_b200.optim import SGD
    from kubeml_b200.parallel.arena import ArenaLayout, WeightArena

    def mk():
        torch.manual_seed(1)
        return torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3), torch.nn.BatchNorm2d(8), torch.nn.Flatten(),
                                   torch.nn.Linear(8 * 36, 10))
    m1, m2 = mk(), mk()
    lay = ArenaLayout(m2)
    assert lay.n_i64 == 1 and lay.n_param % 64 == 0 and lay.n_f32 >= lay.n_param
    idx = lay.index("job1")
    assert idx[0]["key"] == "job1:0.weight" and lay.index("job1", 3)[0]["key"] == "job1:0.weight/3"
    arena = WeightArena(lay, torch.device("cpu"))
    arena.bind(m2)
    o1 = torch.optim.SGD(m1.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    o2 = SGD(m2.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    o2.attach_arena(arena)
    x, y = torch.randn(4, 3, 8, 8), torch.randint(0, 10, (4,))
    for _ in range(3):
        for m, o in ((m1, o1), (m2, o2)):
            o.zero_grad()
            torch.nn.functional.cross_entropy(m(x), y).backward()
            o.step()
    assert o2._flat_views() is not None
    for a, b in zip(m1.state_dict().values(), m2.state_dict().values()):
        torch.testing.assert_close(a, b)
    assert int(arena.counters[0]) == 3
    o2.reset_state()
    assert float(arena.state1.abs().max()) == 0.0


def test_auto_splits_never_leaves_an_empty_split():
    for tiles in (1, 2, 9, 16, 64, 144):
        for chunks in range(1, 600, 7):
            s = nv.auto_splits(tiles, chunks)
            per = -(-chunks // s)
            assert s >= 1 and (s - 1) * per < chunks


def test_vgg11_program_matches_torchvision(fp32_activations):
    import torchvision
    from kubeml_b200.models import vgg11
    torch.manual_seed(0)
    net = vgg11(num_classes=10, dropout=0.0)
    net._ensure_arena()
    with torch.no_grad():
        for p in net.parameters():
            if p.dim() >= 2:
                p.copy_(p.bfloat16().float())
            else:
                p.uniform_(-0.1, 0.1)               # non-zero biases so their path is exercised
    net._arena.refresh_bf16()
    ref = torchvision.models.vgg11(num_classes=10, dropout=0.0)
    assert list(net.state_dict().keys()) == list(ref.state_dict().keys())
    assert all(a.shape == b.shape for a, b in zip(net.state_dict().values(), ref.state_dict().values()))
    ref.load_state_dict(net.state_dict())
    net.train(); ref.train()
    x = torch.randn(4, 3, 32, 32)
    y = torch.randint(0, 10, (4,))
    out = net(x)
    ops.cross_entropy(out, y).backward()
    out_r = ref(x)
    torch.nn.functional.cross_entropy(out_r, y).backward()
    torch.testing.assert_close(out, out_r, rtol=1e-4, atol=1e-4)
    for n, p in ref.named_parameters():
        g = net._arena.grad(n)
        scale = p.grad.abs().max().item() + 1e-12
        assert (g - p.grad).abs().max().item() / scale < 1e-3, n


def test_vgg_dropout_train_and_eval(fp32_activations):
    from kubeml_b200.models import vgg11
    torch.manual_seed(0)
    net = vgg11(num_classes=10, dropout=0.5)
    net._ensure_arena()
    x = torch.randn(2, 3, 32, 32)
    net.train()
    out = net(x)
    a = out.detach().clone()                          # the logits buffer is reused by the next step
    out.float().sum().backward()
    assert int(net._counter) == 1                     # bumped once per step
    b = net(x)
    assert not torch.equal(a, b)                      # fresh mask
    net.eval()
    with torch.no_grad():
        assert torch.equal(net(x), net(x))


def test_plan_cluster_prefers_short_k_loops():
    """Tile width / split choice of the cluster split-K GEMMs: ResNet-34 shapes keep their tuned
    choice, long-K VGG shapes get wider tiles or two CTAs per SM, nothing exceeds the cluster size."""
    cases = {  # (rows, channels, K chunks) -> (block_n, splits)
        (8192, 64, 9): (64, 2), (2048, 128, 18): (64, 4), (512, 256, 36): (64, 8), (128, 512, 72): (64, 8),
        (2048, 512, 72): (128, 4), (8192, 256, 36): (128, 2), (131072, 64, 9): (64, 1),
    }
    for (M, n, chunks), want in cases.items():
        mt = (M + 127) // 128
        got = nv.plan_cluster(mt, n, chunks, nv.pick_block_n(mt, n))
        assert got == want, ((M, n, chunks), got)
        assert 1 <= got[1] <= 8
    for tiles in (1, 9, 64, 144):
        for chunks in range(1, 400, 13):
            s = nv.auto_splits(tiles, chunks, min_chunks=1, budget=296)
            per = -(-chunks // s)
            assert s >= 1 and (s - 1) * per < chunks and tiles * s <= max(296, tiles)


def test_lenet_program_matches_reference_network(fp32_activations):
    """The padded-operand LeNet program against the reference's torch module (function_lenet.LeNet)."""
    from kubeml_b200.functions.function_lenet import LeNet
    from kubeml_b200.models import lenet
    torch.manual_seed(0)
    net = lenet()
    net._ensure_arena()
    ref = LeNet()
    assert list(net.state_dict().keys()) == list(ref.state_dict().keys())
    assert all(a.shape == b.shape for a, b in zip(net.state_dict().values(), ref.state_dict().values()))
    with torch.no_grad():
        for p in net.parameters():
            p.copy_((p * 2).bfloat16().float())          # bf16-exact values, large enough for live ReLUs
    ref.load_state_dict(net.state_dict())
    net.train(); ref.train()
    x = torch.randn(16, 1, 28, 28).bfloat16().float()
    y = torch.randint(0, 10, (16,))
    out = net(x)
    ops.cross_entropy(out, y).backward()
    out_r = ref(x)
    torch.nn.functional.cross_entropy(out_r, y).backward()
    torch.testing.assert_close(out, out_r, rtol=1e-4, atol=1e-4)
    for n, p in ref.named_parameters():
        g = net._arena.grad(n)
        scale = p.grad.abs().max().item() + 1e-12
        assert (g - p.grad).abs().max().item() / scale < 1e-3, n
    # trains under the stock optimizer the reference's function uses
    opt = torch.optim.SGD(net.parameters(), lr=0.05, momentum=0.9)
    l0 = None
    for _ in range(5):
        opt.zero_grad()
        loss = ops.cross_entropy(net(x), y)
        loss.backward()
        opt.step()
        l0 = l0 if l0 is not None else float(loss)
    assert float(loss) < l0


def test_cifar_resnet20_program_matches_torch_module(fp32_activations):
    """Native CIFAR ResNet (option-A shortcuts, 16/32/64 channels) against the stock-module version."""
    from kubeml_b200.models import cifar_native, cifar_resnet
    torch.manual_seed(0)
    net = cifar_native.resnet20(num_classes=10)
    net._ensure_arena()
    with torch.no_grad():
        for p in net.parameters():
            if p.dim() >= 2:
                p.copy_(p.bfloat16().float())
    net._arena.refresh_bf16()
    ref = cifar_resnet.resnet20(num_classes=10)
    assert list(net.state_dict().keys()) == list(ref.state_dict().keys())
    ref.load_state_dict(net.state_dict())
    net.train(); ref.train()
    x = torch.randn(4, 3, 32, 32)
    y = torch.randint(0, 10, (4,))
    out = net(x)
    ops.cross_entropy(out, y).backward()
    out_r = ref(x)
    torch.nn.functional.cross_entropy(out_r, y).backward()
    torch.testing.assert_close(out, out_r, rtol=1e-4, atol=1e-4)
    for n, p in ref.named_parameters():
        g = net._arena.grad(n)
        scale = p.grad.abs().max().item() + 1e-12
        assert (g - p.grad).abs().max().item() / scale < 2e-3, n
    assert int(net.layer3[2].bn2.num_batches_tracked) == 1
