#!/usr/bin/env python
"""Final-loss parity: native bf16 kernels vs the stock fp32 path, same recipe, same data, same init.

BASELINE.json asks for "final-loss parity" with the reference next to the images/s number.  The
reference's recipe is ml/experiments/kubeml/function_resnet34.py:62-80 (torchvision resnet34(), SGD
lr 0.1 wd 1e-4 no momentum, F.cross_entropy, K-AVG every K=8 batches of 128 per worker).  Two arms
train that recipe for the same number of steps on the same synthetic-but-learnable data (CIFAR-shaped
uint8 images, labels from a fixed random "teacher" so there is something to learn; uniform-random labels
would only measure memorisation noise), from the same initial state_dict, without augmentation:

  --impl native     this repository: `TrainJob` over `SpmdRuntime`, hand-written bf16 kernels with fp32
                    master weights, fused merge over peer memory
  --impl ref-torch  stock torchvision modules in fp32 (cuDNN/cuBLAS; TF32 off), torch.optim.SGD, the
                    reference's contiguous doc sharding, NCCL all_reduce / N every K steps, optimizer reset

Both print one JSON line with the per-epoch training loss (mean over workers of their mean batch loss, the
reference's History.train_loss, ml/pkg/train/util.go:82-95), validation loss and accuracy.

    python bench/loss_parity.py --impl native --gpus 1 --epochs 16 --steps-per-epoch 32
    python -m torch.distributed.run --nproc-per-node 8 ... bench/loss_parity.py --impl ref-torch --gpus 8 ...
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
BATCH, K_AVG, LR, WD = 128, 8, 0.1, 1e-4
MEAN = (0.485, 0.456, 0.406)
STD = (0.229, 0.224, 0.225)


def make_data(n_train: int, n_test: int, seed: int = 7):
    """uint8 [n,32,32,3] images with spatial structure and labels = argmax of a fixed random linear
    teacher over the 4x4-pooled image: 10 classes, learnable, deterministic."""
    import numpy as np
    rng = np.random.default_rng(seed)
    n = n_train + n_test
    low = rng.normal(size=(n, 8, 8, 3)).astype(np.float32)
    img = np.repeat(np.repeat(low, 4, axis=1), 4, axis=2) + 0.5 * rng.normal(size=(n, 32, 32, 3)).astype(np.float32)
    x = np.clip(128 + 48 * img, 0, 255).astype(np.uint8)
    pooled = x.reshape(n, 8, 4, 8, 4, 3).astype(np.float32).mean(axis=(2, 4)).reshape(n, -1)
    pooled = (pooled - pooled.mean(0)) / (pooled.std(0) + 1e-6)
    teacher = np.random.default_rng(seed + 1).normal(size=(pooled.shape[1], 10)).astype(np.float32)
    y = (pooled @ teacher).argmax(1).astype(np.int64)
    return x[:n_train], y[:n_train], x[n_train:], y[n_train:]


def init_state_path(home: str) -> str:
    """One initial state_dict for both arms (the native network carries torchvision's names/shapes)."""
    import torch
    path = os.path.join(home, "init_resnet34.pt")
    if not os.path.isfile(path):
        from torchvision.models.resnet import resnet34
        torch.manual_seed(1234)
        torch.save(resnet34().state_dict(), path)
    return path


def relaunch(args) -> int:
    port = os.environ.get("MASTER_PORT", "29533")
    return subprocess.call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                            f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port", port,
                            os.path.abspath(__file__)] + sys.argv[1:])


def run_native(args, home: str) -> dict:
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    os.environ["KUBEML_HOME"] = home
    from kubeml_b200.api.types import JobInfo, JobState, TrainOptions, TrainRequest, TrainTask
    from kubeml_b200.runtime.spmd import SpmdRuntime
    from kubeml_b200.storage.store import ShardStore
    from kubeml_b200.train.history import HistoryStore
    from kubeml_b200.train.job import TrainJob
    N = args.gpus
    store = ShardStore(os.path.join(home, "datasets"))
    name = f"parity_n{N}_s{args.steps_per_epoch}"
    if rank == 0 and not store.exists(name):
        xtr, ytr, xte, yte = make_data(N * args.steps_per_epoch * BATCH, 2048)
        store.create(name, xtr, ytr, xte, yte)
    init = init_state_path(home) if rank == 0 else os.path.join(home, "init_resnet34.pt")
    rt = SpmdRuntime(shard_store=store)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    if rank != 0:
        rt.serve()
        rt.close()
        return {}
    req = TrainRequest(model_type="resnet34", batch_size=BATCH, epochs=args.epochs, dataset=name, lr=LR,
                       function_name="resnet34",
                       options=TrainOptions(default_parallelism=N, static_parallelism=True, validate_every=1,
                                            k=K_AVG, goal_accuracy=101.0))
    task = TrainTask(request=req, job=JobInfo(id="parity", state=JobState(parallelism=N)))
    inv = rt.invoker()
    inv.env = {"KUBEML_RESNET_DATASET": name, "KUBEML_RESNET_IMPL": "native", "KUBEML_RESNET_NOAUG": "1",
               "KUBEML_RESNET_INIT": init}
    job = TrainJob(task, inv, rt.tensor_store, history_store=HistoryStore(os.path.join(home, "history")),
                   device_mode=True)
    job.train()
    rt.stop()
    rt.close()
    if job.exit_err is not None:
        raise SystemExit(f"parity job failed: {job.exit_err}")
    h = job.history
    return {"train_loss": [round(v, 5) for v in h.train_loss], "validation_loss": [round(v, 5) for v in h.validation_loss],
            "accuracy": [round(v, 3) for v in h.accuracy], "dtype": "bf16 operands, fp32 accumulate and master weights"}


def run_ref_torch(args, home: str) -> dict:
    import numpy as np
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    from collections import defaultdict
    from torchvision.models.resnet import resnet34
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.allow_tf32 = False            # the fp32 the reference's torch 1.7.1 computed in
    torch.backends.cuda.matmul.allow_tf32 = False
    N = args.gpus
    xtr, ytr, xte, yte = make_data(N * args.steps_per_epoch * BATCH, 2048)
    if rank == 0:
        init_state_path(home)
    if world > 1:
        dist.barrier()
    net = resnet34()
    net.load_state_dict(torch.load(os.path.join(home, "init_resnet34.pt"), map_location="cpu"))
    net = net.to(dev)
    mean = torch.tensor(MEAN, device=dev).view(1, 3, 1, 1) * 255.0
    inv = 1.0 / (torch.tensor(STD, device=dev).view(1, 3, 1, 1) * 255.0)

    def prep(xb):
        return (torch.as_tensor(xb, device=dev).permute(0, 3, 1, 2).float() - mean) * inv

    # contiguous shard per worker (python/kubeml/kubeml/util.py:46-56): rank r takes the r-th slice
    per = args.steps_per_epoch * BATCH
    xs, ys = xtr[rank * per:(rank + 1) * per], torch.as_tensor(ytr[rank * per:(rank + 1) * per], device=dev)
    vper = len(xte) // N
    xv, yv = xte[rank * vper:(rank + 1) * vper], torch.as_tensor(yte[rank * vper:(rank + 1) * vper], device=dev)

    def merge():
        with torch.no_grad():
            for t in net.state_dict().values():
                if world > 1:
                    dist.all_reduce(t)
                if t.is_floating_point():
                    t.div_(world)
                else:
                    t.copy_(torch.div(t, world, rounding_mode="trunc"))

    out = {"train_loss": [], "validation_loss": [], "accuracy": []}
    for epoch in range(args.epochs):
        net.train()
        opt = torch.optim.SGD(net.parameters(), lr=LR, weight_decay=WD)     # fresh per epoch like configure_optimizers
        tot = torch.zeros((), device=dev)
        for s in range(args.steps_per_epoch):
            opt.zero_grad()
            loss = F.cross_entropy(net(prep(xs[s * BATCH:(s + 1) * BATCH])), ys[s * BATCH:(s + 1) * BATCH])
            loss.backward()
            opt.step()
            tot += loss.detach()
            if (s + 1) % K_AVG == 0 or s + 1 == args.steps_per_epoch:
                merge()
                opt.state = defaultdict(dict)
        tl = tot / args.steps_per_epoch
        net.eval()
        with torch.no_grad():
            vl, acc, nb = torch.zeros((), device=dev), torch.zeros((), device=dev), 0
            for s in range(0, vper, BATCH):
                o = net(prep(xv[s:s + BATCH]))
                vl += F.cross_entropy(o, yv[s:s + BATCH])
                acc += (o.argmax(1) == yv[s:s + BATCH]).float().sum() * (100.0 / BATCH)
                nb += 1
            stats = torch.stack([tl, vl / nb, acc / nb])
        if world > 1:
            dist.all_reduce(stats)
            stats /= world
        tl_, vl_, ac_ = (float(v) for v in stats.tolist())
        out["train_loss"].append(round(tl_, 5)); out["validation_loss"].append(round(vl_, 5)); out["accuracy"].append(round(ac_, 3))
    out["dtype"] = "fp32 (TF32 off)"
    if world > 1:
        dist.destroy_process_group()
    return out if rank == 0 else {}


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="native", choices=["native", "ref-torch"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--epochs", type=int, default=16)
    ap.add_argument("--steps-per-epoch", type=int, default=32)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            return relaunch(args)
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    home = os.path.join(tempfile.gettempdir(), f"kubeml_parity_{os.getuid()}")
    os.makedirs(home, exist_ok=True)
    res = run_native(args, home) if args.impl == "native" else run_ref_torch(args, home)
    if res:
        res.update(impl=args.impl, n_gpus=args.gpus, epochs=args.epochs, steps_per_epoch=args.steps_per_epoch,
                   total_steps=args.epochs * args.steps_per_epoch, batch=BATCH, k=K_AVG, lr=LR, weight_decay=WD,
                   model="resnet34 (1000 logits) on 32x32x3", data="synthetic, teacher-labelled, no augmentation")
        print(json.dumps(res), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
