import sys, os, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kubeml_b200 import ops
from kubeml_b200.models import resnet18
variant = sys.argv[1]
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
y = torch.randint(0, 10, (32,), device=dev)
if variant == "A":
    lin = torch.nn.Linear(64, 1000).to(dev)
    xin = torch.randn(32, 64, device=dev)
    def step():
        loss = ops.cross_entropy(lin(xin), y); loss.backward(); return loss
else:
    net = resnet18(num_classes=1000).to(dev); net._ensure_arena()
    x = torch.randn(32, 32, 32, 3, device=dev).bfloat16()
    dl = torch.randn(32, 1000, device=dev).bfloat16()
    if variant == "B":
        def step():
            out = net(x); out.float().sum().backward(); return out
    elif variant == "C":
        def step():
            with torch.no_grad():
                out = net._forward_impl(x, train=True); net._backward_impl(dl)
            return out
    elif variant in ("E", "F", "G"):
        import kubeml_b200.models.native as nv
        dummy = torch.zeros(1, device=dev, requires_grad=True)
        zero = torch.zeros(1, device=dev)
        class Fn(torch.autograd.Function):
            @staticmethod
            def forward(ctx, x, anchor):
                ctx.anchor = anchor
                return net._forward_impl(x, train=True)
            @staticmethod
            def backward(ctx, dlog):
                net._backward_impl(dlog.contiguous())
                if variant == "E": return None, zero
                if variant == "F": return None, None
                return None, torch.zeros_like(ctx.anchor)
        anchor = net.fc.bias if variant == "G" else dummy
        def step():
            out = Fn.apply(x, anchor); out.float().sum().backward(); return out
    elif variant == "D":
        def step():
            out = net(x); torch.autograd.backward(out, grad_tensors=dl); return out
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for _ in range(2): step()
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(g):
        step()
    g.replay(); torch.cuda.synchronize()
    print("ok  ", variant, flush=True)
except Exception as e:
    print("FAIL", variant, str(e).splitlines()[0][:150], flush=True)
