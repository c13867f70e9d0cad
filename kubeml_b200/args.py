"""Per-invocation arguments a worker receives (reference `_KubeArgs`,
python/kubeml/kubeml/dataset.py:24-78).  The reference ships them as URL query parameters
(train/function.go:44-68); the same names are kept so `to_query()`/`parse()` round-trip."""
from __future__ import annotations

from typing import Any, Mapping, Optional

from .exceptions import InvalidArgsError


class _KubeArgs:
    def __init__(self, job_id: str, N: int, K: int, task: str, func_id: int, epoch: int,
                 lr: float = 0.0, batch_size: int = 0):
        self._job_id = job_id
        self._N = N
        self._K = K
        self._task = task
        self._func_id = func_id
        self.lr = lr
        self.batch_size = batch_size
        self.epoch = epoch

    @classmethod
    def parse(cls, query: Optional[Mapping[str, Any]] = None) -> "_KubeArgs":
        """Build from a query mapping; with no argument, read the worker's current invocation
        context (the stand-in for flask.request, dataset.py:63-78)."""
        if query is None:
            from .runtime.context import current_invocation
            query = current_invocation().query

        def get(name, typ, default=None):
            v = query.get(name, default)
            if v is None:
                return None
            return typ(v)
        try:
            return cls(job_id=get("jobId", str), N=get("N", int), K=get("K", int),
                       task=get("task", str), func_id=get("funcId", int),
                       epoch=get("epoch", int), lr=get("lr", float, 0.0),
                       batch_size=get("batchSize", int, 0))
        except (ValueError, TypeError) as e:
            raise InvalidArgsError(e)

    def to_query(self) -> dict:
        return {"task": self._task, "jobId": self._job_id, "N": self._N, "K": self._K,
                "funcId": self._func_id, "batchSize": self.batch_size, "lr": self.lr,
                "epoch": self.epoch}

    def __repr__(self) -> str:
        return f"_KubeArgs({self.to_query()})"
