"""Analysis helpers of the reference notebooks (ml/experiments/insights.ipynb cells 3-4, 13-14;
online_learn.ipynb cell 17), as importable functions.

* `tta(acc, times, goal)`           first epoch whose validation accuracy reaches `goal`
* `tta_crossbow(acc, times, goal)`  first epoch whose *median of the last 5 accuracies* reaches it
* `total_time(history)`            `epoch_duration[-1]` (cumulative wall clock incl. validations)
* `KOptimizer`                     the notebook's online-learning prototype: predict training time and
                                   final accuracy from (batch, lr, N, K) with PassiveAggressiveRegressor,
                                   updated after every finished experiment (not wired into the scheduler
                                   there either)
"""
from __future__ import annotations

import math
from statistics import median
from typing import List, Optional, Sequence


def tta(accuracy: Sequence[float], epoch_duration: Sequence[float], goal: float) -> float:
    for a, t in zip(accuracy, epoch_duration):
        if a >= goal:
            return float(t)
    return math.nan


def tta_crossbow(accuracy: Sequence[float], epoch_duration: Sequence[float], goal: float, window: int = 5) -> float:
    for i in range(len(accuracy)):
        lo = max(0, i - window + 1)
        if median(accuracy[lo:i + 1]) >= goal and i + 1 >= min(window, len(accuracy)):
            return float(epoch_duration[i])
    return math.nan


def total_time(epoch_duration: Sequence[float]) -> float:
    return float(epoch_duration[-1]) if len(epoch_duration) else math.nan


def mean_parallelism(parallelism: Sequence[float]) -> float:
    return float(sum(parallelism) / len(parallelism)) if len(parallelism) else math.nan


class KOptimizer:
    """Online regressors time ~ f(batch, lr, N, K) and accuracy ~ g(batch, lr, N, K)."""

    def __init__(self):
        from sklearn.linear_model import PassiveAggressiveRegressor
        # features are already on log scales of similar magnitude; time is learned in log space
        self.time_model = PassiveAggressiveRegressor(C=0.5, epsilon=0.01, random_state=0)
        self.acc_model = PassiveAggressiveRegressor(C=0.5, epsilon=0.1, random_state=0)
        self._n = 0

    @staticmethod
    def features(batch: int, lr: float, n: int, k: int) -> List[float]:
        k_eff = 256.0 if k == -1 else float(k)          # "sparse averaging": one sync per epoch
        return [math.log2(batch) / 8.0, math.log10(lr) / 3.0, float(n) / 8.0, math.log2(k_eff) / 8.0, 1.0]

    def update(self, batch: int, lr: float, n: int, k: int, time_s: float, accuracy: float) -> None:
        xs = [self.features(batch, lr, n, k)]
        self.time_model.partial_fit(xs, [math.log(max(time_s, 1e-6))])
        self.acc_model.partial_fit(xs, [accuracy])
        self._n += 1

    def predict(self, batch: int, lr: float, n: int, k: int):
        if self._n == 0:
            return math.nan, math.nan
        xs = [self.features(batch, lr, n, k)]
        return float(math.exp(self.time_model.predict(xs)[0])), float(self.acc_model.predict(xs)[0])

    def best_k(self, batch: int, lr: float, n: int, candidates=(-1, 32, 16, 8), min_accuracy: Optional[float] = None):
        scored = [(self.predict(batch, lr, n, k), k) for k in candidates]
        ok = [(t, k) for (t, a), k in scored if min_accuracy is None or a >= min_accuracy]
        return min(ok or [(t, k) for (t, a), k in scored])[1]
