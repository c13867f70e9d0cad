"""Experiment harness (reference: ml/experiments/common/experiment.py:19-250, common/utils.py,
train.py, app/time_to_accuracy.py).

`KubemlExperiment.run()` submits a train task, starts system-metrics collection, polls until the
task left the running list, fetches its History and stores everything in one record; `run_grid()`
walks a (batch x K x N) grid, skips the points already on disk (the reference's grid-level resume,
train.py:93-96) and retries transient failures with exponential back-off (common/utils.py:83-120).
The reference shells out to the `kubeml` binary; this harness talks to the same controller through
the typed client (in-process cluster or HTTP), so it needs no subprocess.
"""
from __future__ import annotations

import functools
import itertools
import json
import os
import time
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterable, List, Optional

import numpy as np

from ..api.types import History, TrainOptions, TrainRequest
from ..controller.client import KubemlClient


def retry(exceptions=Exception, total_tries: int = 5, initial_wait: float = 0.5, backoff_factor: float = 2.0):
    """Decorator: retry with exponential back-off (common/utils.py:83-120)."""
    def deco(fn):
        @functools.wraps(fn)
        def wrapper(*a, **k):
            wait = initial_wait
            for attempt in range(1, total_tries + 1):
                try:
                    return fn(*a, **k)
                except exceptions:
                    if attempt == total_tries:
                        raise
                    time.sleep(wait)
                    wait *= backoff_factor
        return wrapper
    return deco


# the reference's grids (train.py:17-61, app/time_to_accuracy.py:79-85)
LENET_GRID = {"batch": [128, 64, 32, 16], "k": [-1, 32, 16, 8], "parallelism": [1, 2, 4, 8]}
RESNET_GRID = {"batch": [256, 128, 64, 32], "k": [-1, 32, 16, 8], "parallelism": [2, 4, 8]}
EPOCHS = 30


@dataclass
class ExperimentRecord:
    title: str
    request: dict
    task_id: str = ""
    history: dict = field(default_factory=dict)
    system_metrics: dict = field(default_factory=dict)
    wall_seconds: float = 0.0
    error: Optional[str] = None

    def to_json(self) -> str:
        return json.dumps(self.__dict__)


class KubemlExperiment:
    def __init__(self, title: str, request: TrainRequest, client: KubemlClient,
                 metrics=None, poll_seconds: float = 1.0):
        self.title, self.request, self.client = title, request, client
        self.metrics = metrics                  # experiments.sysmetrics.MetricsCollector or None
        self.poll_seconds = poll_seconds
        self.network_id: Optional[str] = None
        self.history: Optional[History] = None

    @retry(Exception, total_tries=5, backoff_factor=2, initial_wait=0.5)
    def run_task(self) -> str:
        return self.client.v1().networks().train(self.request)

    @retry(Exception, total_tries=5, backoff_factor=2, initial_wait=0.5)
    def check_if_task_finished(self) -> bool:
        return all(t.job.id != self.network_id for t in self.client.v1().tasks().list())

    @retry(Exception, total_tries=5, backoff_factor=2, initial_wait=0.5)
    def get_model_history(self) -> History:
        return self.client.v1().histories().get(self.network_id)

    def wait_for_task_finished(self, timeout: Optional[float] = None) -> None:
        t0 = time.time()
        time.sleep(min(self.poll_seconds, 0.2))           # let the scheduler create the job
        while not self.check_if_task_finished():
            if timeout is not None and time.time() - t0 > timeout:
                raise TimeoutError(self.network_id)
            time.sleep(self.poll_seconds)

    def run(self, timeout: Optional[float] = None) -> ExperimentRecord:
        rec = ExperimentRecord(self.title, self.request.to_dict())
        t0 = time.time()
        try:
            self.network_id = self.run_task()
            rec.task_id = self.network_id
            if self.metrics is not None:
                self.metrics.start(self.network_id)
            self.wait_for_task_finished(timeout)
            self.history = self.get_model_history()
            rec.history = self.history.to_dict()
        except Exception as e:  # noqa: BLE001 - a failed grid point is recorded, not fatal
            rec.error = str(e)
        finally:
            if self.metrics is not None:
                rec.system_metrics = self.metrics.finish()
            rec.wall_seconds = time.time() - t0
        return rec

    @staticmethod
    def _fake_history(epochs: int = 3) -> History:
        """Dry-run plumbing without a cluster (common/experiment.py:204-216)."""
        from ..api.types import JobHistory
        xs = [float(i) for i in range(1, epochs + 1)]
        return History(id="fake", data=JobHistory(validation_loss=[1 / x for x in xs], accuracy=[10 * x for x in xs],
                                                  train_loss=[1 / x for x in xs], parallelism=[2.0] * epochs,
                                                  epoch_duration=xs))


def grid_points(grid: Dict[str, Iterable]) -> List[dict]:
    keys = list(grid)
    return [dict(zip(keys, vals)) for vals in itertools.product(*(grid[k] for k in keys))]


def point_title(function: str, p: dict) -> str:
    return f"{function}-b{p['batch']}-k{p['k']}-p{p['parallelism']}"


def run_grid(client: KubemlClient, function: str, dataset: str, lr: float, grid: Dict[str, Iterable],
             out_dir: str, epochs: int = EPOCHS, validate_every: int = 1, goal_accuracy: float = 100.0,
             metrics=None, on_result: Optional[Callable[[ExperimentRecord], None]] = None,
             timeout: Optional[float] = None) -> List[ExperimentRecord]:
    """Run every grid point that has no result file yet (resume = recompute the missing points)."""
    os.makedirs(out_dir, exist_ok=True)
    done = []
    for p in grid_points(grid):
        title = point_title(function, p)
        path = os.path.join(out_dir, title + ".json")
        if os.path.isfile(path):
            with open(path) as f:
                d = json.load(f)
            if not d.get("error"):
                done.append(ExperimentRecord(**d))
                continue
        req = TrainRequest(model_type=function, batch_size=p["batch"], epochs=epochs, dataset=dataset, lr=lr,
                           function_name=function,
                           options=TrainOptions(default_parallelism=p["parallelism"], static_parallelism=True,
                                                validate_every=validate_every, k=p["k"],
                                                goal_accuracy=goal_accuracy))
        rec = KubemlExperiment(title, req, client, metrics=metrics).run(timeout)
        with open(path, "w") as f:
            f.write(rec.to_json())
        done.append(rec)
        if on_result is not None:
            on_result(rec)
    return done


def to_dataframe(records: List[ExperimentRecord]):
    """One row per experiment with the columns the reference notebooks use."""
    import pandas as pd
    rows = []
    for r in records:
        opt = r.request.get("options", {})
        h = r.history.get("data", {}) if r.history else {}
        rows.append({"id": r.task_id, "title": r.title, "model_type": r.request.get("model_type"),
                     "batch_size": r.request.get("batch_size"), "epochs": r.request.get("epochs"),
                     "lr": r.request.get("lr"), "k": opt.get("k"), "default_parallelism": opt.get("default_parallelism"),
                     "static_parallelism": opt.get("static_parallelism"),
                     "validation_loss": h.get("validation_loss", []), "accuracy": h.get("accuracy", []),
                     "train_loss": h.get("train_loss", []), "parallelism": h.get("parallelism", []),
                     "epoch_duration": h.get("epoch_duration", []), "error": r.error})
    return pd.DataFrame(rows)


# ---------------------------------------------------------------------------------- the comparison system
class StockTorchExperiment:
    """The harness side of the comparison system: one plain-PyTorch training run, in this process, recorded the
    way a KubeML experiment is (reference: `TensorflowExperiment` runs the Keras LeNet / ResNet-34 baselines
    in-process under MirroredStrategy and keeps their per-epoch metrics plus the cumulative seconds of a
    `TimeHistory` callback, ml/experiments/common/experiment.py:263-337, tflow/time_callback.py:6-12).

    The model is any stock `nn.Module` factory (`function_lenet.LeNet`, `torchvision.models.resnet34`, ...), the
    data a dataset of the shard store, the optimizer SGD with momentum 0.9 and L2 1e-4 (tflow/lenet.py,
    tflow/resnet34.py); under `torch.distributed` (torchrun) the module is wrapped in DistributedDataParallel and
    every rank reads its contiguous share - the gradient all-reduce system KubeML is compared with.  The record's
    `history.data` has the same keys as a KubeML history; `epoch_duration` is per epoch and `times` cumulative."""

    def __init__(self, title: str, model_factory: Callable, dataset: str, batch: int, epochs: int, lr: float,
                 store=None, device: Optional[str] = None, transform: Optional[Callable] = None, metrics=None,
                 momentum: float = 0.9, weight_decay: float = 1e-4):
        self.title, self.model_factory, self.dataset = title, model_factory, dataset
        self.batch, self.epochs, self.lr = int(batch), int(epochs), float(lr)
        self.store, self.device, self.transform, self.metrics = store, device, transform, metrics
        self.momentum, self.weight_decay = momentum, weight_decay
        self.times: List[float] = []

    def run(self) -> ExperimentRecord:
        import torch
        import torch.distributed as dist
        import torch.nn.functional as F
        from ..storage import default_store
        store = self.store or default_store()
        dev = torch.device(self.device or ("cuda" if torch.cuda.is_available() else "cpu"))
        rank, world = (dist.get_rank(), dist.get_world_size()) if dist.is_available() and dist.is_initialized() else (0, 1)
        rec = ExperimentRecord(self.title, {"model_type": self.title.split("-")[0], "batch_size": self.batch,
                                            "epochs": self.epochs, "lr": self.lr, "dataset": self.dataset,
                                            "options": {"default_parallelism": world, "static_parallelism": True, "k": 1}})
        t_start = time.time()
        if self.metrics is not None:
            self.metrics.start(self.title)
        try:
            xtr, ytr = store.arrays(self.dataset, "train")
            xte, yte = store.arrays(self.dataset, "test")
            per = -(-len(xtr) // world)
            lo, hi = rank * per, min((rank + 1) * per, len(xtr))
            prep = self.transform or (lambda x: x.float().div(255.0).unsqueeze(1) if x.dim() == 3
                                      else x.float().div(255.0).permute(0, 3, 1, 2))
            net = self.model_factory().to(dev)
            model = torch.nn.parallel.DistributedDataParallel(net) if world > 1 else net
            opt = torch.optim.SGD(model.parameters(), lr=self.lr, momentum=self.momentum, weight_decay=self.weight_decay)
            hist = {"train_loss": [], "validation_loss": [], "accuracy": [], "parallelism": [], "epoch_duration": []}
            self.times = []
            for _ in range(self.epochs):
                t0 = time.time()
                model.train()
                tot, nb = 0.0, 0
                for i in range(lo, hi, self.batch):
                    x = prep(torch.from_numpy(np.array(xtr[i:min(i + self.batch, hi)])).to(dev))
                    y = torch.from_numpy(np.array(ytr[i:min(i + self.batch, hi)])).to(dev).long().view(-1)
                    opt.zero_grad()
                    loss = F.cross_entropy(model(x), y)
                    loss.backward()
                    opt.step()
                    tot, nb = tot + float(loss.detach()), nb + 1
                model.eval()
                vl, correct, n = 0.0, 0, 0
                with torch.no_grad():
                    for i in range(0, len(xte), self.batch):
                        x = prep(torch.from_numpy(np.array(xte[i:i + self.batch])).to(dev))
                        y = torch.from_numpy(np.array(yte[i:i + self.batch])).to(dev).long().view(-1)
                        out = net(x)
                        vl += float(F.cross_entropy(out, y, reduction="sum"))
                        correct += int((out.argmax(1) == y).sum())
                        n += len(y)
                hist["train_loss"].append(tot / max(nb, 1))
                hist["validation_loss"].append(vl / max(n, 1))
                hist["accuracy"].append(100.0 * correct / max(n, 1))
                hist["parallelism"].append(float(world))
                hist["epoch_duration"].append(time.time() - t0)
                self.times.append(time.time() - t_start)          # TimeHistory: cumulative seconds at each epoch end
            rec.history = {"_id": self.title, "task": rec.request, "data": hist, "times": list(self.times)}
        except Exception as e:  # noqa: BLE001 - recorded like a failed grid point
            rec.error = str(e)
        finally:
            if self.metrics is not None:
                rec.system_metrics = self.metrics.finish()
            rec.wall_seconds = time.time() - t_start
        return rec
