"""Epoch timeline as a Chrome trace (`chrome://tracing`, Perfetto) — the reference has no tracing
beyond a debug log line around the merge (ml/pkg/train/job.go:397,412)."""
from __future__ import annotations

import json
import threading
import time
from contextlib import contextmanager
from typing import List


class ChromeTrace:
    def __init__(self):
        self._ev: List[dict] = []
        self._lock = threading.Lock()
        self._t0 = time.perf_counter()

    def _us(self) -> float:
        return (time.perf_counter() - self._t0) * 1e6

    @contextmanager
    def span(self, name: str, cat: str = "job", tid: int = 0, **args):
        t = self._us()
        try:
            yield
        finally:
            with self._lock:
                self._ev.append({"name": name, "cat": cat, "ph": "X", "ts": t, "dur": self._us() - t,
                                 "pid": 0, "tid": tid, "args": args})

    def instant(self, name: str, cat: str = "job", tid: int = 0, **args) -> None:
        with self._lock:
            self._ev.append({"name": name, "cat": cat, "ph": "i", "ts": self._us(), "pid": 0, "tid": tid,
                             "s": "g", "args": args})

    def device_span(self, name: str, start_us: float, dur_us: float, tid: int, **args) -> None:
        """A span measured with CUDA events (e.g. one merge) placed on a worker's row."""
        with self._lock:
            self._ev.append({"name": name, "cat": "device", "ph": "X", "ts": start_us, "dur": dur_us, "pid": 1,
                             "tid": tid, "args": args})

    def save(self, path: str) -> None:
        with self._lock, open(path, "w") as f:
            json.dump({"traceEvents": self._ev, "displayTimeUnit": "ms"}, f)
