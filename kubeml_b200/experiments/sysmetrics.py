"""System-metrics sidecar (reference: ml/experiments/common/metrics.py:23-160): while an experiment
runs, sample CPU %, memory and per-GPU load / memory every `interval` seconds on a thread, and hand
back the series when it finishes.  The reference samples with psutil + GPUtil behind a Flask app
(`PUT /new/<id>`, `DELETE /finish`); this one samples with psutil + NVML and offers the same two
routes on the shared HTTP plumbing."""
from __future__ import annotations

import threading
import time
from typing import Dict, List, Optional

from ..httpd import Router, Server


def _nvml():
    try:
        import pynvml
        pynvml.nvmlInit()
        return pynvml
    except Exception:  # noqa: BLE001 - no driver on the build box
        return None


class MetricsCollector:
    def __init__(self, interval: float = 2.0):
        self.interval = interval
        self._thread: Optional[threading.Thread] = None
        self._stop = threading.Event()
        self._rows: List[dict] = []
        self.experiment_id: Optional[str] = None

    def _sample(self, nv) -> dict:
        import psutil
        vm = psutil.virtual_memory()
        row = {"time": time.time(), "cpu": psutil.cpu_percent(None), "mem": vm.percent}
        if nv is not None:
            try:
                for i in range(nv.nvmlDeviceGetCount()):
                    h = nv.nvmlDeviceGetHandleByIndex(i)
                    u = nv.nvmlDeviceGetUtilizationRates(h)
                    m = nv.nvmlDeviceGetMemoryInfo(h)
                    row[f"gpu_{i}"] = u.gpu / 100.0
                    row[f"gpu_{i}_mem_used"] = m.used / 2 ** 20
                    row[f"gpu_{i}_mem_free"] = m.free / 2 ** 20
                    row[f"gpu_{i}_sm_mhz"] = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            except Exception:  # noqa: BLE001
                pass
        return row

    def _loop(self):
        nv = _nvml()
        while not self._stop.is_set():
            self._rows.append(self._sample(nv))
            self._stop.wait(self.interval)

    def start(self, experiment_id: str) -> None:
        if self._thread is not None:
            self.finish()
        self.experiment_id = experiment_id
        self._rows = []
        self._stop.clear()
        self._thread = threading.Thread(target=self._loop, daemon=True, name="sysmetrics")
        self._thread.start()

    def finish(self) -> Dict[str, list]:
        if self._thread is None:
            return {}
        self._stop.set()
        self._thread.join(timeout=5)
        self._thread = None
        keys = sorted({k for r in self._rows for k in r})
        return {"id": self.experiment_id, **{k: [r.get(k) for r in self._rows] for k in keys}}

    # the reference's two routes
    def router(self) -> Router:
        r = Router()
        r.add("PUT", "/new/{id}", lambda q: (self.start(q.params["id"]), {"started": q.params["id"]})[1])
        r.add("DELETE", "/finish", lambda q: self.finish())
        r.add("GET", "/health", lambda q: (200, None))
        return r

    def serve(self, host: str = "127.0.0.1", port: int = 0) -> Server:
        return Server(self.router(), host, port).start()
