"""`python -m kubeml_b200.experiments ...` - the experiment drivers of the reference as one command
(ml/experiments/train.py:17-96 runs the LeNet / ResNet grids batch x K x N with static parallelism and validation
every epoch; ml/experiments/app/time_to_accuracy.py:60-110 the same grids against a goal accuracy; both resume by
recomputing only the grid points without a result file; the notebooks then read time-to-accuracy off the histories).

    python -m kubeml_b200.experiments --function lenet --dataset mnist --lr 0.01 --epochs 30 \
        --batch 128 64 32 16 --k -1 32 16 8 --parallelism 1 2 4 8 --goal 99 --out results/lenet [--url URL | --local]

Talks to a running control plane (`--url`, default $KUBEML_CONTROLLER_URL) or starts one in this process (`--local`);
samples CPU / memory / GPU load next to every run (`--sysmetrics`); prints one line per grid point with total time,
final accuracy and the two time-to-accuracy definitions, and leaves `<out>/<title>.json` per point plus
`<out>/summary.json`.  `--grid lenet|resnet` takes the reference's grids as they are.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from typing import List, Optional

from . import analysis, harness


def _num(v: float) -> Optional[float]:
    return None if v != v else round(float(v), 3)


def summarise(records: List[harness.ExperimentRecord], goal: float) -> List[dict]:
    rows = []
    for r in records:
        h = r.history.get("data", {}) if r.history else {}
        acc, dur = h.get("accuracy", []), h.get("epoch_duration", [])
        n = min(len(acc), len(dur))          # the final validation has no epoch of its own
        opt = r.request.get("options", {})
        rows.append({"title": r.title, "task": r.task_id, "error": r.error, "batch": r.request.get("batch_size"),
                     "k": opt.get("k"), "parallelism": opt.get("default_parallelism"),
                     "total_time_s": _num(analysis.total_time(dur)) if dur else None,
                     "final_accuracy": round(acc[-1], 2) if acc else None,
                     "tta_s": _num(analysis.tta(acc[:n], dur[:n], goal)) if n else None,
                     "tta_crossbow_s": _num(analysis.tta_crossbow(acc[:n], dur[:n], goal)) if n else None})
    return rows


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(prog="python -m kubeml_b200.experiments", description=__doc__.split("\n\n")[0])
    ap.add_argument("--function", "-f", default=None)
    ap.add_argument("--dataset", "-d", default=None)
    ap.add_argument("--lr", type=float, default=None)
    ap.add_argument("--epochs", "-e", type=int, default=harness.EPOCHS)
    ap.add_argument("--grid", choices=["lenet", "resnet"], default=None, help="the reference's grid for that network")
    ap.add_argument("--batch", type=int, nargs="+", default=None)
    ap.add_argument("--k", type=int, nargs="+", default=None)
    ap.add_argument("--parallelism", type=int, nargs="+", default=None)
    ap.add_argument("--validate-every", type=int, default=1)
    ap.add_argument("--goal", type=float, default=100.0, help="goal accuracy (stops a run early, and the TTA threshold)")
    ap.add_argument("--out", required=True)
    ap.add_argument("--url", default=None)
    ap.add_argument("--local", action="store_true", help="run the control plane inside this process")
    ap.add_argument("--sysmetrics", action="store_true")
    ap.add_argument("--timeout", type=float, default=None, help="seconds per grid point")
    args = ap.parse_args(argv)

    presets = {"lenet": ("lenet", "mnist", 0.01, harness.LENET_GRID), "resnet": ("resnet34", "cifar10", 0.1, harness.RESNET_GRID)}
    fn, ds, lr, grid = presets[args.grid] if args.grid else (None, None, None, {"batch": [64], "k": [-1], "parallelism": [2]})
    grid = {"batch": args.batch or grid["batch"], "k": args.k or grid["k"], "parallelism": args.parallelism or grid["parallelism"]}
    fn, ds, lr = args.function or fn, args.dataset or ds, args.lr if args.lr is not None else lr
    if not fn or not ds or lr is None:
        ap.error("--function, --dataset and --lr are required (or --grid lenet|resnet)")

    from ..controller.client import make_kubeml_client
    cluster = None
    if args.local:
        from ..controller.cluster import Cluster
        cluster = Cluster()
        client = make_kubeml_client(cluster)
    else:
        client = make_kubeml_client(url=args.url)
    metrics = None
    if args.sysmetrics:
        from .sysmetrics import MetricsCollector
        metrics = MetricsCollector()
    try:
        def show(rec):
            row = summarise([rec], args.goal)[0]
            print(json.dumps(row), flush=True)
        recs = harness.run_grid(client, fn, ds, lr, grid, args.out, epochs=args.epochs, validate_every=args.validate_every,
                                goal_accuracy=args.goal, metrics=metrics, on_result=show, timeout=args.timeout)
    finally:
        if cluster is not None:
            cluster.shutdown()
    rows = summarise(recs, args.goal)
    with open(os.path.join(args.out, "summary.json"), "w") as f:
        json.dump({"function": fn, "dataset": ds, "lr": lr, "epochs": args.epochs, "goal": args.goal, "rows": rows}, f, indent=1)
    failed = [r["title"] for r in rows if r["error"]]
    print(f"{len(rows) - len(failed)} of {len(rows)} grid points have a result in {args.out}"
          + (f"; failed: {', '.join(failed)}" if failed else ""), flush=True)
    return 1 if failed else 0


if __name__ == "__main__":
    sys.exit(main())
