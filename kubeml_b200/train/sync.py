"""How a worker exchanges weights with the rest of the job.

Two engines implement the same five verbs used by `KubeModel`:

* `HostSync`   — the reference's data path with RedisAI replaced by the in-process TensorStore:
                 publish the whole state_dict per round, block on the job's barrier, reload the
                 merged model (network.py:208-224, 395-461).  CPU jobs and the oracle use it.
* `DeviceSync` — kubeml_b200.parallel.device_sync: parameters *are* slices of a flat HBM arena
                 mapped into every peer GPU, the merge is one sm_100a kernel that reads peers,
                 averages and writes every replica, and the barrier is a flag in peer memory.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import numpy as np
import torch

from ..exceptions import MergeError, StorageError
from ..model.store import TensorStore, weight_key


class HostSync:
    name = "host"

    def __init__(self, store: TensorStore, barrier: Optional[Callable[[str, int], None]]):
        self.store = store
        self.barrier = barrier

    # -- init task ---------------------------------------------------------------------------
    def publish_reference(self, job_id: str, network: torch.nn.Module) -> List[str]:
        sd = network.state_dict()
        self.store.multi_set({weight_key(job_id, n): _np(t) for n, t in sd.items()})
        return list(sd.keys())

    # -- per epoch -------------------------------------------------------------------------
    def attach(self, kube_model) -> None:
        """Nothing to bind on the host path."""

    # -- per round ---------------------------------------------------------------------------
    def load_reference(self, job_id: str, network: torch.nn.Module) -> None:
        names = list(network.state_dict().keys())
        try:
            arrs = self.store.multi_get(weight_key(job_id, n) for n in names)
        except KeyError as e:
            raise StorageError(e)
        network.load_state_dict({n: torch.from_numpy(np.array(a, copy=True))
                                 for n, a in zip(names, arrs)})

    def publish_replica(self, job_id: str, func_id: int, network: torch.nn.Module) -> None:
        with torch.no_grad():
            self.store.multi_set({weight_key(job_id, n, func_id): _np(t)
                                  for n, t in network.state_dict().items()})

    def arrive(self, job_id: str, func_id: int, round_idx: int) -> None:
        if self.barrier is None:
            raise MergeError(RuntimeError("no train job to synchronise with"))
        self.barrier(job_id, func_id)

    def end_epoch(self, job_id: str, func_id: int, rounds_done: int) -> None:
        """The last round's merge is triggered by the function's return (function.go:180-190)."""

    def drain_losses(self, loss):
        return float(loss)


def _np(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().numpy()
