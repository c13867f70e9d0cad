"""`bench.py --impl ref-torch`: the reference's *algorithm* with stock libraries only (SURVEY 6.3 #1,
BASELINE.md "ref-torch"): one process per GPU, torchvision resnet34 (cuDNN/cuBLAS, eager, fp32 — what
torch 1.7.1 would run for the reference), `F.cross_entropy`, `torch.optim.SGD`, K local steps, then
`dist.all_reduce` (NCCL) of every state_dict tensor divided by N (integer division for int64) and an
optimizer-state reset.  No kernel, model or engine of kubeml_b200 is used.  Inputs are staged on the
device (this is the kernel-side comparison; the reference arm covers the end-to-end path).
`--impl ddp` is the gradient-all-reduce data-parallel baseline the reference compares itself against
(its TensorFlow MirroredStrategy runs, ml/experiments/tflow/resnet34.py:56; TensorFlow is not
installed here, torch DDP is the same strategy)."""
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
    from torchvision.models.resnet import resnet34

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
    bf16 = os.environ.get("KUBEML_REF_TORCH_BF16") == "1"     # stronger stock baseline: channels-last + bf16 autocast
    net = resnet34().to(dev)
    if bf16:
        net = net.to(memory_format=torch.channels_last)
    if ddp and world > 1:
        net = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local])
    opt = torch.optim.SGD(net.parameters(), lr=0.1, weight_decay=1e-4)
    g = torch.Generator(device=dev).manual_seed(rank)
    total = args.warmup + args.steps
    x = torch.randn(total, BATCH, 3, 32, 32, device=dev, generator=g)
    y = torch.randint(0, 10, (total, BATCH), device=dev, generator=g)

    def merge():
        from collections import defaultdict
        with torch.no_grad():
            for t in net.state_dict().values():
                if world > 1:
                    dist.all_reduce(t)
                if t.is_floating_point():
                    t.div_(world)
                else:
                    t.copy_(torch.div(t, world, rounding_mode="trunc"))
        opt.state = defaultdict(dict)

    def step(i):
        opt.zero_grad()
        if bf16:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = net(x[i].contiguous(memory_format=torch.channels_last))
            loss = F.cross_entropy(out.float(), y[i])
        else:
            loss = F.cross_entropy(net(x[i]), y[i])
        loss.backward()
        opt.step()
        if not ddp and (i + 1) % K_AVG == 0:
            merge()
        return loss

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.warmup, total):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 1e3], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        value = world * args.steps * BATCH / float(t.item())
        print(json.dumps({"metric": "images/sec ResNet-34 " + ("DDP gradient all-reduce" if ddp else "K-AVG (K=8, batch 128/worker)"),
                          "impl": "ddp" if ddp else "ref-torch", "value": round(value, 1), "unit": "images/s",
                          "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                          "ms_per_step": round(1e3 * float(t.item()) / args.steps, 3), "higher_is_better": True,
                          "scaling": "weak", "vs_baseline": round(value / BASELINE_IMG_S, 2), "dtype": "bf16" if bf16 else "fp32",
                          "data": "synthetic (device-resident)",
                          "config": {"model": "torchvision resnet34() 1000 logits on 32x32x3", "global_batch": BATCH * world,
                                     "seq_len": None, "parallelism": ("ddp" if ddp else "kavg-dp") + str(world), "k": K_AVG}}),
              flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0
