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


class CollectiveSync:
    """K-AVG through `torch.distributed` collectives (gloo on CPU, NCCL on GPUs): the stock-library
    form of the merge — `all_reduce(SUM)` of every state_dict tensor, then divide by the number of
    contributors, integer division for int64 (SURVEY 6.3 "ref-torch").  It serves as the NCCL
    baseline the fused merge kernel is measured against and as the multi-process CPU test path.
    Restricted to static parallelism N == world size."""
    name = "collective"

    def __init__(self, store: TensorStore, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.store = store
        self.group = group
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self._model = None
        self._max_rounds = 0
        self._my_rounds = 0
        self._done = 0
        self._fresh = True
        self.merge_seconds = 0.0

    def publish_reference(self, job_id, network):
        sd = network.state_dict()
        self.store.multi_set({weight_key(job_id, n): _np(t) for n, t in sd.items()})
        return list(sd.keys())

    def attach(self, kube_model) -> None:
        from ..util import round_schedule
        self._model = kube_model
        a = kube_model.args
        if a._N != self.world:
            raise MergeError(RuntimeError("CollectiveSync needs parallelism == world size"))
        if a._task == "train":
            sched = round_schedule(kube_model._dataset.num_docs, a._N, a._K, a.batch_size)
            self._max_rounds = len(sched)
            self._my_rounds = sum(1 for s in sched if self.rank in s)
        self._done = 0
        self._fresh = True
        self.merge_seconds = 0.0

    def load_reference(self, job_id, network):
        if not self._fresh:
            return          # the previous merge already left the averaged model in `network`
        self._fresh = False
        if self.rank == 0:
            names = list(network.state_dict().keys())
            arrs = self.store.multi_get(weight_key(job_id, n) for n in names)
            network.load_state_dict({n: torch.from_numpy(np.array(a, copy=True))
                                     for n, a in zip(names, arrs)})
        if self.world > 1:
            for t in network.state_dict().values():
                self.dist.broadcast(t, src=0, group=self.group)

    def publish_replica(self, job_id, func_id, network):
        pass

    def _allreduce_average(self, network, contributes: bool) -> None:
        import time
        dev = next(network.parameters()).device
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        cnt = torch.tensor([1.0 if contributes else 0.0], device=dev)
        if self.world > 1:
            self.dist.all_reduce(cnt, group=self.group)
        n = int(cnt.item())
        with torch.no_grad():
            for t in network.state_dict().values():
                buf = t if contributes else torch.zeros_like(t)
                if self.world > 1:
                    self.dist.all_reduce(buf, group=self.group)
                if contributes and n > 0:
                    if t.is_floating_point():
                        t.div_(n)
                    else:
                        t.copy_(torch.div(t, n, rounding_mode="trunc"))
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        self.merge_seconds += time.perf_counter() - t0

    def arrive(self, job_id, func_id, round_idx):
        self._allreduce_average(self._model._network, True)
        self._done += 1

    def end_epoch(self, job_id, func_id, rounds_done):
        net = self._model._network
        if rounds_done > 0:
            self._allreduce_average(net, True)
            self._done += 1
        while self._done < self._max_rounds:      # keep the collectives aligned with busier ranks
            self._allreduce_average(net, False)
            self._done += 1
        # every rank keeps the model it holds as its local copy of the reference
        self.store.multi_set({weight_key(job_id, n): _np(t) for n, t in net.state_dict().items()})

    def drain_losses(self, loss):
        return float(loss)
