"""Host tensor store: the in-process stand-in for RedisAI.

Key layout is the reference's (ml/pkg/model/utils.go:140-158, python/kubeml/kubeml/network.py:456-458):
reference model `"{jobId}:{state_dict_key}"`, worker replica `"{jobId}:{state_dict_key}/{funcId}"`.
Values are typed row-major arrays (float32 weights/buffers, int64 `num_batches_tracked`), copied
on set like a Redis write so later in-place training cannot alias a published replica.

The device-resident equivalent (same names, slices of a flat HBM arena) lives in
kubeml_b200.parallel.arena; this host store serves CPU jobs, the correctness oracle and
checkpoints.
"""
from __future__ import annotations

import fnmatch
import threading
from typing import Dict, Iterable, List, Mapping, Optional

import numpy as np


def weight_key(job_id: str, layer_name: str, func_id: int = -1) -> str:
    """`func_id < 0` addresses the reference/init model (utils.go:148-156)."""
    return f"{job_id}:{layer_name}/{func_id}" if func_id >= 0 else f"{job_id}:{layer_name}"


class TensorStore:
    def __init__(self):
        self._d: Dict[str, np.ndarray] = {}
        self._lock = threading.Lock()

    def tensorset(self, key: str, value) -> None:
        arr = np.array(value, copy=True, order="C")
        with self._lock:
            self._d[key] = arr

    def multi_set(self, items: Mapping[str, np.ndarray]) -> None:
        """All-or-nothing publish, like the MULTI/EXEC block of Model.Save (model.go:135-161)."""
        staged = {k: np.array(v, copy=True, order="C") for k, v in items.items()}
        with self._lock:
            self._d.update(staged)

    def tensorget(self, key: str) -> np.ndarray:
        with self._lock:
            try:
                return self._d[key]
            except KeyError:
                raise KeyError(f"tensor {key!r} not in store") from None

    def multi_get(self, keys: Iterable[str]) -> List[np.ndarray]:
        with self._lock:
            return [self._d[k] for k in keys]

    def exists(self, key: str) -> bool:
        with self._lock:
            return key in self._d

    def keys(self, pattern: str = "*") -> List[str]:
        with self._lock:
            return [k for k in self._d if fnmatch.fnmatchcase(k, pattern)]

    def delete(self, keys: Iterable[str]) -> int:
        n = 0
        with self._lock:
            for k in list(keys):
                if self._d.pop(k, None) is not None:
                    n += 1
        return n

    def nbytes(self) -> int:
        with self._lock:
            return sum(v.nbytes for v in self._d.values())


_default: Optional[TensorStore] = None


def default_tensor_store() -> TensorStore:
    global _default
    if _default is None:
        _default = TensorStore()
    return _default
