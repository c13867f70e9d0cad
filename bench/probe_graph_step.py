"""Bisect CUDA-graph capture of a native training step (bring-up diagnostic)."""
import sys, os, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import ops
from kubeml_b200.models import resnet18
from kubeml_b200.optim import SGD
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
net = resnet18(num_classes=1000).to(dev)
net._ensure_arena()
opt = SGD(net.parameters(), lr=0.1, weight_decay=1e-4)
opt.attach_arena(net._arena)
x = torch.randn(32, 32, 32, 3, device=dev).bfloat16()
y = torch.randint(0, 10, (32,), device=dev)

def stage(level):
    opt.zero_grad() if level >= 4 else None
    out = net(x)
    if level >= 2:
        loss = ops.cross_entropy(out, y)
        if level >= 3:
            loss.backward()
        if level >= 4:
            opt.step()
        return loss
    return out

side = torch.cuda.Stream()
for level in (3, 4):
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(2):
            stage(level)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g):
            r = stage(level)
        g.replay(); torch.cuda.synchronize()
        import time
        t0 = time.perf_counter()
        for _ in range(20): g.replay()
        torch.cuda.synchronize()
        print(f"ok   level {level}  replay {1e3*(time.perf_counter()-t0)/20:.3f} ms", flush=True)
    except Exception as e:
        print(f"FAIL level {level}: {str(e).splitlines()[0][:200]}", flush=True)
        traceback.print_exc(limit=12)
        break
