"""Throughput-based dynamic parallelism (reference: ml/pkg/scheduler/policy.go:9-102).

Per job the policy remembers the *reference* epoch time.  First sight: hand out the default
parallelism and create the task.  Second sight (no reference yet): +1.  Afterwards compare the
last epoch's elapsed time with the reference:

    elapsed <= ref * 1.05  -> +1 worker, new reference      (scaling up did not hurt)
    elapsed >= ref * 1.20  -> -1 worker, new reference      (it got clearly slower)
    otherwise              -> keep N, keep the reference

The reference never clamps, so N can reach 0 or exceed the hardware; this
implementation clamps to [1, max_parallelism] and takes its lock around every cache write
(the reference writes `timeCache` unlocked in three branches).
"""
from __future__ import annotations

import enum
import threading
from typing import Dict, Protocol, Tuple

from ..api.const import DEFAULT_PARALLELISM, MAX_GPUS
from ..api.types import TrainTask

SCALE_DOWN_THRESHOLD = 1.2
SCALE_UP_THRESHOLD = 1.05


class TaskOperation(enum.Enum):
    CREATE = "create"
    UPDATE = "update"


class SchedulerPolicy(Protocol):
    def calculate_parallelism(self, task: TrainTask) -> Tuple[int, TaskOperation]: ...
    def task_finished(self, task_id: str) -> None: ...


class ThroughputBasedPolicy:
    def __init__(self, max_parallelism: int = MAX_GPUS, min_parallelism: int = 1):
        self.max_parallelism = max_parallelism
        self.min_parallelism = min_parallelism
        self._time_cache: Dict[str, float] = {}
        self._mu = threading.Lock()

    def _clamp(self, n: int) -> int:
        return max(self.min_parallelism, min(self.max_parallelism, n))

    def calculate_parallelism(self, task: TrainTask) -> Tuple[int, TaskOperation]:
        jid = task.job.id
        elapsed = task.job.state.elapsed_time
        current = task.job.state.parallelism
        with self._mu:
            if jid not in self._time_cache:
                self._time_cache[jid] = 0.0
                n = task.request.options.default_parallelism or DEFAULT_PARALLELISM
                return self._clamp(n), TaskOperation.CREATE
            prev = self._time_cache[jid]
            if prev == 0:
                self._time_cache[jid] = elapsed
                return self._clamp(current + 1), TaskOperation.UPDATE
            if elapsed <= prev * SCALE_UP_THRESHOLD:
                self._time_cache[jid] = elapsed
                return self._clamp(current + 1), TaskOperation.UPDATE
            if elapsed >= prev * SCALE_DOWN_THRESHOLD:
                self._time_cache[jid] = elapsed
                return self._clamp(current - 1), TaskOperation.UPDATE
            return self._clamp(current), TaskOperation.UPDATE

    def task_finished(self, task_id: str) -> None:
        with self._mu:
            self._time_cache.pop(task_id, None)

    def reference_time(self, task_id: str):
        with self._mu:
            return self._time_cache.get(task_id)
