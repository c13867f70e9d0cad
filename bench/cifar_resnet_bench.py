"""ResNet-32 / CIFAR-10 step time: native kernels (eager and CUDA-graph replay) against the stock torch
modules under channels-last bf16 autocast, batch 128, SGD momentum 0.9."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import ops
from kubeml_b200.models import cifar_native, cifar_resnet
from kubeml_b200.optim import SGD
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
x = torch.randn(128, 32, 32, 3, device=dev).bfloat16()
y = torch.randint(0, 10, (128,), device=dev)


def timeit(fn, n=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


net = cifar_native.resnet32().to(dev); net._ensure_arena()
opt = SGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4); opt.attach_arena(net._arena)
net.train()
def step():
    opt.zero_grad(); loss = ops.cross_entropy(net(x), y); loss.backward(); opt.step(); return loss
t_eager = timeit(step)
side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    step(); g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        step()
torch.cuda.current_stream().wait_stream(side)
t_graph = timeit(g.replay)
ref = cifar_resnet.resnet32().to(dev).to(memory_format=torch.channels_last); ref.train()
ropt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
xr = x.permute(0, 3, 1, 2).float()
def rstep():
    ropt.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = ref(xr)
    loss = torch.nn.functional.cross_entropy(out.float(), y); loss.backward(); ropt.step(); return loss
t_ref = timeit(rstep)
print(f"resnet32 batch 128: native eager {t_eager:.3f} ms, native graph {t_graph:.3f} ms ({128 / t_graph:.1f} K img/s), "
      f"stock torch bf16 eager {t_ref:.3f} ms ({128 / t_ref:.1f} K img/s)")
