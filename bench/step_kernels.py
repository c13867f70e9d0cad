"""Eager training steps of a native model for a per-kernel time list:
  python bench/step_kernels.py vgg16 3                                  (must exit 0 first)
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/k.csv \
      python bench/step_kernels.py vgg16 3
  python bench/step_kernels.py --summarize gpurun_out/k.csv 3           (last step only)"""
import csv, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def summarize(path: str, steps: int) -> None:
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("==")) if len(r) > 5]
    hdr = rows[0]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    gs = hdr.index("Grid Size")
    ks = [(r[kn].split("(")[0].replace("void ", "").replace("<unnamed>::", ""), r[gs],
           float(r[mv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[mu], 1e-3)) for r in rows[1:]]
    per = len(ks) // steps
    last = ks[-per:]
    total = sum(t for _, _, t in last)
    print(f"{per} kernels per step, {total:.1f} us of kernel time in the last step\n")
    agg = {}
    for n, g, t in last:
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1; a[1] += t
    print("| kernel | launches | total us | share |\n|---|---|---|---|")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{n[:90]}` | {c} | {t:.1f} | {100 * t / total:.1f}% |")
    print("\nin launch order (kernels >= 20 us):")
    for i, (n, g, t) in enumerate(last):
        if t >= 20:
            print(f"  #{i:3d} {t:8.1f} us  grid {g:>14s}  {n[:100]}")


if __name__ == "__main__":
    if sys.argv[1] == "--summarize":
        summarize(sys.argv[2], int(sys.argv[3]))
        sys.exit(0)
    import torch
    from kubeml_b200 import models, ops
    from kubeml_b200.optim import SGD, Adam
    name, steps = sys.argv[1], int(sys.argv[2])
    dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
    if name.startswith("cifar_"):
        net = getattr(models.cifar_native, name[6:])(num_classes=10).to(dev)
    else:
        net = getattr(models, name)(num_classes=1000).to(dev)
    net._ensure_arena()
    opt = Adam(net.parameters(), lr=1e-4) if name.startswith("vgg") else SGD(net.parameters(), lr=0.1, weight_decay=1e-4)
    opt.attach_arena(net._arena)
    net.train()
    x = torch.randn(128, 32, 32, 3, device=dev).bfloat16()
    y = torch.randint(0, 100, (128,), device=dev)
    for _ in range(steps):
        opt.zero_grad()
        loss = ops.cross_entropy(net(x), y)
        loss.backward()
        opt.step()
    torch.cuda.synchronize()
    print("ok", float(loss))
