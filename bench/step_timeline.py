"""Where a captured ResNet-34 training step spends its time, by stage, without a profiler:
    KUBEML_STAMPS=1 python bench/step_timeline.py [replays]
The native network writes the GPU's nanosecond timer at stage boundaries (`kml_stamp`, 13 one-thread kernels per
step); the step is captured in a CUDA graph like the training job does and replayed; the last replay's stamps
are printed as per-stage durations (forward, backward, optimizer), next to the whole-step time from CUDA events."""
import os, sys
os.environ["KUBEML_STAMPS"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import models, ops
from kubeml_b200.optim import SGD

replays = int(sys.argv[1]) if len(sys.argv) > 1 else 50
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
net = models.resnet34(num_classes=1000).to(dev)
net._ensure_arena()
opt = SGD(net.parameters(), lr=0.1, weight_decay=1e-4)
opt.attach_arena(net._arena)
net.train()
x = torch.randn(128, 32, 32, 3, device=dev).bfloat16()
y = torch.randint(0, 10, (128,), device=dev)
end = torch.zeros(1, dtype=torch.int64, device=dev)


def step():
    opt.zero_grad()
    loss = ops.cross_entropy(net(x), y)
    loss.backward()
    opt.step()
    ops.nn.stamp(end, 0)
    return loss


for _ in range(3):
    step()
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    loss = step()
for _ in range(10):
    g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(replays):
    g.replay()
e1.record()
torch.cuda.synchronize()
st = net.stamps.cpu().tolist()
names = net.STAMP_NAMES
t_end = int(end.item())
print(f"step (CUDA events, {replays} replays): {e0.elapsed_time(e1) / replays * 1e3:.1f} us")
prev = st[0]
rows = []
for i in range(1, len(names)):
    rows.append((names[i], (st[i] - prev) / 1e3))
    prev = st[i]
rows.append(("optimizer (+ side-stream join)", (t_end - prev) / 1e3))
tot = (t_end - st[0]) / 1e3
for n, d in rows:
    print(f"  {n:32s} {d:8.1f} us  {100 * d / tot:5.1f} %")
print(f"  {'start -> end of step':32s} {tot:8.1f} us")
