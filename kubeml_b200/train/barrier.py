"""Merge barrier with a shrinking participant set.

Reference mechanics (ml/pkg/train/job.go:368-442, api.go:100-126, function.go:180-190): a
merger goroutine waits on a WaitGroup sized to the still-running functions; each function either
*arrives* (POST /next, blocks for the verdict) or *returns* (its last round); both add their
replica to the accumulator; when everyone is accounted for the accumulator is divided by the
number of contributions, saved, and the arrivers are released; the WaitGroup is re-armed with
the functions that have not returned yet.

This implementation has no merger thread: whoever completes the round performs the merge
("last one in closes the door"), which removes the channel/WaitGroup reuse hazards listed in
SURVEY 5.2.  One deliberate fix: a function that *failed* is released from the
barrier but contributes no replica, so the divisor counts only replicas actually summed.
"""
from __future__ import annotations

import threading
import time
from typing import Callable, List, Optional

from ..exceptions import MergeError
from ..model.model import Model, ParallelSGD


class MergeBarrier:
    def __init__(self, model: Model, optimizer: ParallelSGD, parallelism: int,
                 on_merge: Optional[Callable[[List[int], float], None]] = None,
                 timeout: Optional[float] = None):
        self.model = model
        self.optimizer = optimizer
        self.running = parallelism          # functions that have not returned yet
        self.pending = parallelism          # of those, how many still owe this round an arrival
        self.contributors: List[int] = []   # replicas summed in this round, in arrival order
        self.generation = 0
        self.failed: Optional[BaseException] = None
        self.done = threading.Event()       # set when every function returned and the last merge ran
        self.rounds = 0
        self.merge_seconds = 0.0
        self._cv = threading.Condition()
        self._on_merge = on_merge
        self._timeout = timeout
        self.model.clear()

    # -- called by a worker that finished a non-final round (blocks) ---------------------------
    def arrive(self, func_id: int) -> None:
        self._account(func_id, contributes=True, returning=False)

    # -- called when a function returned (ok) or failed (not ok); never blocks on peers ---------
    def finished(self, func_id: int, ok: bool) -> None:
        self._account(func_id, contributes=ok, returning=True)

    def _account(self, func_id: int, contributes: bool, returning: bool) -> None:
        if self.failed is not None:
            if returning:
                with self._cv:
                    self.running -= 1
                return
            raise MergeError(self.failed)
        if contributes:
            try:
                self.model.update(func_id)       # outside the barrier lock: fetches overlap
            except Exception as e:               # noqa: BLE001  (missing replica: model.go:261-266)
                contributes = False
                if not returning:
                    with self._cv:
                        self.failed = e
                        self._cv.notify_all()
                    raise MergeError(e)
        with self._cv:
            if contributes:
                self.contributors.append(func_id)
            if returning:
                self.running -= 1
            self.pending -= 1
            gen = self.generation
            if self.pending == 0:
                self._merge_locked()
            elif not returning:
                deadline = None if self._timeout is None else time.monotonic() + self._timeout
                while self.generation == gen and self.failed is None:
                    remaining = None if deadline is None else deadline - time.monotonic()
                    if remaining is not None and remaining <= 0:
                        self.failed = TimeoutError("merge barrier timed out")
                        self._cv.notify_all()
                        break
                    self._cv.wait(remaining)
            if not returning and self.failed is not None:
                raise MergeError(self.failed)

    def _merge_locked(self) -> None:
        t0 = time.perf_counter()
        try:
            if not self.contributors:
                raise RuntimeError("no functions returned for merging")
            self.optimizer.average(self.model, len(self.contributors))
            self.model.save()
        except Exception as e:  # noqa: BLE001
            self.failed = e
        dt = time.perf_counter() - t0
        self.merge_seconds += dt
        self.rounds += 1
        if self._on_merge is not None and self.failed is None:
            self._on_merge(list(self.contributors), dt)
        self.contributors = []
        self.model.clear()
        self.pending = self.running
        self.generation += 1
        if self.running == 0 or self.failed is not None:
            self.done.set()
        self._cv.notify_all()
