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
