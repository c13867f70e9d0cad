"""FIFO of train tasks awaiting a scheduling decision (reference: ml/pkg/scheduler/queue.go:15-83,
a mutex-guarded slice polled every 10 ms).  This one blocks on a condition variable instead of
polling."""
from __future__ import annotations

import collections
import threading
from typing import Deque, Optional

from ..api.types import TrainTask


class SchedulerQueue:
    def __init__(self):
        self._q: Deque[TrainTask] = collections.deque()
        self._cv = threading.Condition()
        self._closed = False

    def push(self, task: TrainTask) -> None:
        with self._cv:
            self._q.append(task)
            self._cv.notify()

    def pop(self, timeout: Optional[float] = None) -> Optional[TrainTask]:
        with self._cv:
            if not self._q and not self._closed:
                self._cv.wait(timeout)
            return self._q.popleft() if self._q else None

    def close(self) -> None:
        with self._cv:
            self._closed = True
            self._cv.notify_all()

    def __len__(self) -> int:
        with self._cv:
            return len(self._q)
