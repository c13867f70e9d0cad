"""Prometheus-style gauges with the reference's metric names (ml/pkg/ps/metrics.go:33-133):

    kubeml_job_validation_loss{jobid}      kubeml_job_validation_accuracy{jobid}
    kubeml_job_train_loss{jobid}           kubeml_job_parallelism{jobid}
    kubeml_job_epoch_duration_seconds{jobid}
    kubeml_job_running_total{type="train"|"inference"}

plus two B200-side extensions: kubeml_job_merge_seconds{jobid} (device-timed merge latency per
epoch) and kubeml_job_images_per_second{jobid}.  `render()` emits the text exposition format so a
stock Prometheus can scrape `/metrics` (ps/parameter_server.go:57-66) and the bundled Grafana
dashboard (deploy/grafana/KubeML.json) keeps working.
"""
from __future__ import annotations

import threading
from typing import Dict, Tuple

from ..api.types import MetricUpdate

_JOB_GAUGES = {
    "kubeml_job_validation_loss": ("Validation loss values of the train job", "validation_loss"),
    "kubeml_job_validation_accuracy": ("Validation accuracy of the train job", "accuracy"),
    "kubeml_job_train_loss": ("Train loss values of the train job", "train_loss"),
    "kubeml_job_parallelism": ("Parallelism of the train job", "parallelism"),
    "kubeml_job_epoch_duration_seconds": ("Epoch duration of the train job", "epoch_duration"),
    "kubeml_job_merge_seconds": ("Device-timed K-AVG merge seconds in the last epoch", "merge_seconds"),
    "kubeml_job_images_per_second": ("Training throughput of the last epoch", "images_per_second"),
}


class MetricsRegistry:
    def __init__(self):
        self._mu = threading.Lock()
        self._jobs: Dict[str, Dict[str, float]] = {}
        self._running: Dict[str, int] = {"train": 0, "inference": 0}

    def update(self, job_id: str, m: MetricUpdate) -> None:
        with self._mu:
            self._jobs[job_id] = {name: float(getattr(m, attr))
                                  for name, (_, attr) in _JOB_GAUGES.items()}

    def clear(self, job_id: str) -> None:
        """metrics.go:100-106: drop every series of a finished job."""
        with self._mu:
            self._jobs.pop(job_id, None)

    def task_started(self, kind: str) -> None:
        with self._mu:
            self._running[kind] = self._running.get(kind, 0) + 1

    def task_finished(self, kind: str) -> None:
        with self._mu:
            self._running[kind] = max(0, self._running.get(kind, 0) - 1)

    def get(self, job_id: str) -> Dict[str, float]:
        with self._mu:
            return dict(self._jobs.get(job_id, {}))

    def running(self, kind: str) -> int:
        with self._mu:
            return self._running.get(kind, 0)

    def render(self) -> str:
        with self._mu:
            lines = []
            for name, (help_, _) in _JOB_GAUGES.items():
                lines.append(f"# HELP {name} {help_}")
                lines.append(f"# TYPE {name} gauge")
                for jid, vals in sorted(self._jobs.items()):
                    lines.append(f'{name}{{jobid="{jid}"}} {vals[name]:.9g}')
            lines.append("# HELP kubeml_job_running_total Number of tasks currently running")
            lines.append("# TYPE kubeml_job_running_total gauge")
            for kind, v in sorted(self._running.items()):
                lines.append(f'kubeml_job_running_total{{type="{kind}"}} {v}')
            return "\n".join(lines) + "\n"
