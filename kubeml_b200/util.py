"""Pure helpers shared by workers and the control plane.

Parity: python/kubeml/kubeml/util.py:10-81 (`STORAGE_SUBSET_SIZE`, `get_gpu`, `split_minibatches`,
`get_subset_period`, `is_optimizable`) and ml/pkg/util/utils.go:10-50 (`IsDebugEnv`,
`LimitParallelism`, `FindFreePort`).
"""
from __future__ import annotations

import math
import os
import secrets
import socket
from typing import List

from .api.const import STORAGE_SUBSET_SIZE

__all__ = ["STORAGE_SUBSET_SIZE", "get_gpu", "split_minibatches", "get_subset_period",
           "is_optimizable", "is_debug_env", "limit_parallelism", "find_free_port",
           "create_job_id", "rounds_for", "round_schedule"]


def get_gpu(func_id: int) -> int:
    """Round-robin worker -> GPU, sticky through the GPU_ID env var (util.py:13-34): a warm
    worker process keeps the device it first ran on."""
    try:   # workers that are threads of one process carry their device in the invocation
        from .runtime.context import current_invocation, has_invocation
        if has_invocation():
            ws = current_invocation().worker_state
            if ws is not None and ws.get("gpu_id") is not None:
                return int(ws["gpu_id"])
    except ImportError:
        pass
    pinned = os.environ.get("GPU_ID")
    if pinned is not None:
        return int(pinned)
    import torch
    n = max(torch.cuda.device_count(), 1)
    gpu = func_id % n
    os.environ["GPU_ID"] = str(gpu)
    return gpu


def is_optimizable(layer) -> bool:
    """util.py:37-43 (unused there too): conv and linear layers carry optimisable weights."""
    t = str(type(layer)).lower()
    return "conv" in t or "linear" in t


def split_minibatches(a: range, n: int) -> List[range]:
    """Balanced contiguous split of the subset ids over n workers; the first `len(a) % n`
    workers get one extra (util.py:46-56)."""
    if n <= 0:
        raise ValueError("parallelism must be positive")
    k, m = divmod(len(a), n)
    return [a[i * k + min(i, m):(i + 1) * k + min(i + 1, m)] for i in range(n)]


def get_subset_period(K: int, batch_size: int, assigned_subsets: range) -> int:
    """Subsets consumed per sync round: everything when K == -1, else ceil(B*K/64)
    (util.py:59-81)."""
    if K == -1:
        return len(assigned_subsets)
    return int(math.ceil((batch_size * K) / STORAGE_SUBSET_SIZE))


def rounds_for(num_docs: int, n: int, func_id: int, K: int, batch_size: int) -> range:
    """Round start ids of worker `func_id` (network.py:263-272)."""
    assigned = split_minibatches(range(num_docs), n)[func_id]
    period = get_subset_period(K, batch_size, assigned)
    if period <= 0 or len(assigned) == 0:
        return range(0)
    return range(assigned.start, assigned.stop, period)


def round_schedule(num_docs: int, n: int, K: int, batch_size: int) -> List[List[int]]:
    """For every barrier round of an epoch, the sorted list of workers that publish a replica in
    it.  This is the shrinking participant set of mergeModel (job.go:368-442): a worker with
    fewer rounds drops out of later rounds and the divisor shrinks with it.  Being
    a pure function of (D, N, K, B) lets every GPU worker derive the same schedule locally, so
    the device-side barrier needs no host round-trip."""
    counts = [len(rounds_for(num_docs, n, r, K, batch_size)) for r in range(n)]
    total = max(counts) if counts else 0
    return [[r for r in range(n) if counts[r] > i] for i in range(total)]


def is_debug_env() -> bool:
    """utils.go:26-37."""
    return bool(os.environ.get("DEBUG_ENV", ""))


def limit_parallelism() -> bool:
    """utils.go:40-50: LIMIT_PARALLELISM freezes N even when the scheduler answers otherwise."""
    v = os.environ.get("LIMIT_PARALLELISM", "")
    return v.lower() in ("1", "true", "yes")


def find_free_port() -> int:
    """utils.go:10-23."""
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def create_job_id() -> str:
    """8 hex characters, like uuid.New().String()[:8] (scheduler/util.go:8-10)."""
    return secrets.token_hex(4)
