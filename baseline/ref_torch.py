"""`bench.py --impl ref-torch|ddp`: the reference's *algorithm* with stock libraries only (SURVEY 6.3 #1,
BASELINE.md "ref-torch").  One process per GPU, torchvision model (cuDNN/cuBLAS), `F.cross_entropy`,
`torch.optim.SGD` / `Adam`, K local steps, then an NCCL all-reduce of the state_dict divided by N (integer
division for int64) and an optimizer-state reset.  No kernel, model or engine of kubeml_b200 is used.

`--ref-mode` picks how hard the stock path is pushed, so that "hand-written kernels win" can be told apart
from "CUDA graphs win":
  fp32-eager   what torch 1.7.1 would run for the reference (default)
  bf16-eager   channels-last + bf16 autocast
  bf16-graph   the same step captured in a `torch.cuda.graph` (static inputs, capturable optimizer) -
               the strongest stock cuDNN/cuBLAS configuration on this box
  fp32-graph   fp32 (TF32 allowed, torch's default for cuDNN) under a CUDA graph
The merge all-reduces one flat buffer (the state_dict's float tensors concatenated) - what a competent
user of the stock stack would write - unless KUBEML_REF_TORCH_PER_TENSOR=1 (one all_reduce per tensor, the
literal translation of the reference's per-layer loop).
`--impl ddp` is the gradient-all-reduce data-parallel system the reference compares itself against (its
TensorFlow MirroredStrategy runs, ml/experiments/tflow/resnet34.py:32-79; TensorFlow is not installed here,
torch DDP is the same strategy).
`value` is measured with device-resident inputs; `e2e` repeats the run with every step's uint8 batch copied
host->device from pinned memory and the loss read back once per K steps (what the own arm's e2e does)."""
from __future__ import annotations

import json
import os
import sys

BATCH, K_AVG = 128, 8
BASELINE_IMG_S = 3200.0


def run_ref_torch(args, ddp: bool = False) -> int:
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    from collections import defaultdict

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            import subprocess
            root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
            return subprocess.call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                                    f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port",
                                    os.environ.get("MASTER_PORT", "29521"), os.path.join(root, "bench.py")] + sys.argv[1:])
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    mode = getattr(args, "ref_mode", None) or ("bf16-eager" if os.environ.get("KUBEML_REF_TORCH_BF16") == "1" else "fp32-eager")
    if ddp and mode.endswith("graph"):
        mode = mode.replace("graph", "eager")       # DDP's bucketed all-reduce hooks are not captured here
    bf16, graph = mode.startswith("bf16"), mode.endswith("graph")
    vgg = getattr(args, "model", "resnet34") == "vgg16"
    if vgg:
        from torchvision.models import vgg16
        net = vgg16().to(dev)
    else:
        from torchvision.models.resnet import resnet34
        net = resnet34().to(dev)
    if bf16:
        net = net.to(memory_format=torch.channels_last)
    if ddp and world > 1:
        net = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local])
    if vgg:
        opt = torch.optim.Adam(net.parameters(), lr=torch.tensor(1e-4, device=dev) if graph else 1e-4, capturable=graph)
    else:
        opt = torch.optim.SGD(net.parameters(), lr=0.1, weight_decay=1e-4)
    g = torch.Generator(device=dev).manual_seed(rank)
    total = args.warmup + args.steps
    ncls = 100 if vgg else 10
    x_u8 = torch.randint(0, 256, (total, BATCH, 32, 32, 3), device=dev, generator=g, dtype=torch.uint8)
    y_all = torch.randint(0, ncls, (total, BATCH), device=dev, generator=g)
    x_host = x_u8.cpu().pin_memory()
    y_host = y_all.cpu().pin_memory()
    mean = torch.tensor((0.485, 0.456, 0.406), device=dev).view(1, 3, 1, 1) * 255.0
    inv = 1.0 / (torch.tensor((0.229, 0.224, 0.225), device=dev).view(1, 3, 1, 1) * 255.0)

    per_tensor = os.environ.get("KUBEML_REF_TORCH_PER_TENSOR") == "1"
    module = net.module if hasattr(net, "module") else net
    sd = list(module.state_dict().values())
    floats = [t for t in sd if t.is_floating_point()]
    ints = [t for t in sd if not t.is_floating_point()]

    def merge():
        with torch.no_grad():
            if per_tensor:
                for t in sd:
                    if world > 1:
                        dist.all_reduce(t)
                    if t.is_floating_point():
                        t.div_(world)
                    else:
                        t.copy_(torch.div(t, world, rounding_mode="trunc"))
            else:
                flat = torch.cat([t.reshape(-1) for t in floats])
                if world > 1:
                    dist.all_reduce(flat)
                flat.div_(world)
                torch._foreach_copy_(floats, [c.view_as(t) for c, t in zip(flat.split([t.numel() for t in floats]), floats)])
                if ints:
                    fi = torch.stack([t.reshape(()) for t in ints])
                    if world > 1:
                        dist.all_reduce(fi)
                    fi = torch.div(fi, world, rounding_mode="trunc")
                    torch._foreach_copy_(ints, list(fi.unbind(0)))
        if graph:
            for st in opt.state.values():               # keep the captured state tensors, zero them in place
                for v in st.values():
                    if torch.is_tensor(v):
                        v.zero_()
        else:
            opt.state = defaultdict(dict)

    sx = torch.empty((BATCH, 32, 32, 3), dtype=torch.uint8, device=dev)
    sy = torch.empty((BATCH,), dtype=torch.int64, device=dev)
    sloss = torch.zeros((), device=dev)

    def compute():
        xin = (sx.permute(0, 3, 1, 2).float() - mean) * inv
        opt.zero_grad(set_to_none=False) if graph else opt.zero_grad()
        if bf16:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = net(xin.contiguous(memory_format=torch.channels_last))
            loss = F.cross_entropy(out.float(), sy)
        else:
            loss = F.cross_entropy(net(xin), sy)
        loss.backward()
        opt.step()
        sloss.copy_(loss.detach())

    cgraph = None

    def step(i, e2e):
        nonlocal cgraph
        if e2e:
            sx.copy_(x_host[i], non_blocking=True)
            sy.copy_(y_host[i], non_blocking=True)
        else:
            sx.copy_(x_u8[i]); sy.copy_(y_all[i])
        if graph and cgraph is not None:
            cgraph.replay()
        else:
            compute()
        if not ddp and (i + 1) % K_AVG == 0:
            merge()
            if e2e:
                sloss.item()

    for i in range(min(3, args.warmup)):
        step(i, False)
    if graph:
        torch.cuda.synchronize()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            compute()                                   # side-stream warm-up as the capture docs ask
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        cgraph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(cgraph):
            compute()
    for i in range(min(3, args.warmup), args.warmup):
        step(i, False)

    def timed(e2e: bool) -> float:
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.warmup, total):
            step(i, e2e)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 1e3], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    t_e2e = timed(True)
    t_dev = timed(False)
    if rank == 0:
        value = world * args.steps * BATCH / t_dev
        e2e = world * args.steps * BATCH / t_e2e
        mname = "VGG-16" if vgg else "ResNet-34"
        print(json.dumps({"metric": f"images/sec {mname} " + ("DDP gradient all-reduce" if ddp else "K-AVG (K=8, batch 128/worker)"),
                          "impl": "ddp" if ddp else "ref-torch", "mode": mode, "value": round(value, 1), "unit": "images/s",
                          "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                          "ms_per_step": round(1e3 * t_dev / args.steps, 3), "higher_is_better": True,
                          "scaling": "weak", "vs_baseline": None if vgg else round(value / BASELINE_IMG_S, 2),
                          "dtype": "bf16 autocast" if bf16 else "fp32",
                          "data": "synthetic",
                          "e2e": {"value": round(e2e, 1), "unit": "images/s", "ms_per_step": round(1e3 * t_e2e / args.steps, 3),
                                  "h2d_bytes_per_step": BATCH * (32 * 32 * 3 + 8), "d2h_bytes_per_step": 4.0 / K_AVG},
                          "config": {"model": f"torchvision {'vgg16' if vgg else 'resnet34'}() 1000 logits on 32x32x3",
                                     "global_batch": BATCH * world, "seq_len": None,
                                     "parallelism": ("ddp" if ddp else "kavg-dp") + str(world), "k": K_AVG,
                                     "optimizer": "Adam lr 1e-4" if vgg else "SGD lr 0.1 wd 1e-4",
                                     "merge": "per-tensor all_reduce" if per_tensor else "one flat all_reduce"}}),
              flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0
