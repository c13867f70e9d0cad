"""CPU reference model and K-AVG averaging — the correctness oracle ("ref-cpu").

Behavioural contract distilled from ml/pkg/model/model.go and parallelSGD.go:

* `build()`   fetch `{job}:{layer}` for every layer name returned by init (model.go:76-114);
* `update(f)` fetch `{job}:{layer}/{f}` and, under one lock, *sum* it into the accumulator; the
              first arrival seeds it (model.go:249-302);
* `ParallelSGD.average(m, n)` divide float tensors by float32(n) and int64 tensors by int64(n)
              with integer (truncating) division (parallelSGD.go:26-54);
* `save()`    publish the accumulator as the new reference model atomically (model.go:135-161);
* `clear()`   drop the accumulator before the next round (model.go:117-120).

The GPU merge kernel (csrc/merge.cu) is tested against this file.
"""
from __future__ import annotations

import logging
import threading
from dataclasses import dataclass
from typing import Dict, Iterable, List, Optional

import numpy as np

from .store import TensorStore, weight_key


@dataclass
class Layer:
    name: str
    dtype: str          # "FLOAT" | "INT64" (RedisAI type tags, model/utils.go:95-98)
    weights: np.ndarray

    @staticmethod
    def of(name: str, arr: np.ndarray) -> "Layer":
        if np.issubdtype(arr.dtype, np.integer):
            return Layer(name, "INT64", np.array(arr, dtype=np.int64, copy=True))
        return Layer(name, "FLOAT", np.array(arr, dtype=np.float32, copy=True))


class Model:
    def __init__(self, job_id: str, layer_names: Iterable[str], store: TensorStore,
                 name: str = "", logger: Optional[logging.Logger] = None):
        self.job_id = job_id
        self.name = name
        self.layer_names: List[str] = list(layer_names)
        self.store = store
        self.state_dict: Dict[str, Layer] = {}
        self._mu = threading.Lock()
        self.log = logger or logging.getLogger("kubeml.model")

    def build(self) -> None:
        arrs = self.store.multi_get(weight_key(self.job_id, n) for n in self.layer_names)
        with self._mu:
            self.state_dict = {n: Layer.of(n, a) for n, a in zip(self.layer_names, arrs)}

    def clear(self) -> None:
        with self._mu:
            self.state_dict = {}

    def update(self, func_id: int) -> None:
        """Add worker `func_id`'s published replica into the accumulator.  Fetches overlap
        between callers; the additions are serialised by the model lock like model.go:273-296,
        which also fixes the summation order to arrival order."""
        arrs = self.store.multi_get(weight_key(self.job_id, n, func_id) for n in self.layer_names)
        with self._mu:
            for n, a in zip(self.layer_names, arrs):
                cur = self.state_dict.get(n)
                if cur is None:
                    self.state_dict[n] = Layer.of(n, a)
                elif cur.dtype == "INT64":
                    cur.weights += a.astype(np.int64, copy=False)
                else:
                    np.add(cur.weights, a.astype(np.float32, copy=False), out=cur.weights)

    def save(self) -> None:
        with self._mu:
            self.store.multi_set({weight_key(self.job_id, n): l.weights
                                  for n, l in self.state_dict.items()})

    def num_parameters(self) -> int:
        with self._mu:
            return int(sum(l.weights.size for l in self.state_dict.values()))

    def summary(self) -> str:
        with self._mu:
            lines = [f"Model {self.name or self.job_id}: {len(self.state_dict)} tensors"]
            for n, l in self.state_dict.items():
                lines.append(f"  {n:<48s} {l.dtype:<6s} {tuple(l.weights.shape)}")
            lines.append(f"  total elements: {sum(l.weights.size for l in self.state_dict.values())}")
        text = "\n".join(lines)
        self.log.debug(text)
        return text


class ParallelSGD:
    """parallelSGD.go:26-54.  Stateless; kept as a class so other merge rules can slot in."""

    def average(self, m: Model, num: int) -> None:
        if num <= 0:
            raise ValueError("no functions returned for merging")
        with m._mu:
            for layer in m.state_dict.values():
                if layer.dtype == "INT64":
                    # Go's integer division truncates toward zero (parallelSGD.go:42-43)
                    q = np.abs(layer.weights) // np.int64(num)
                    layer.weights = (np.sign(layer.weights) * q).astype(np.int64)
                else:
                    np.divide(layer.weights, np.float32(num), out=layer.weights)


def kavg_oracle(replicas: List[Dict[str, np.ndarray]]) -> Dict[str, np.ndarray]:
    """One-call form of update*N + average, for tests: rank-order fp32 sum, divide by N,
    int64 truncating division."""
    store = TensorStore()
    names = list(replicas[0].keys())
    for r, sd in enumerate(replicas):
        store.multi_set({weight_key("o", n, r): v for n, v in sd.items()})
    m = Model("o", names, store)
    for r in range(len(replicas)):
        m.update(r)
    ParallelSGD().average(m, len(replicas))
    return {n: m.state_dict[n].weights for n in names}
