"""What a worker needs from its surroundings, bundled so functions stay stateless."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable, Optional

from ..model.store import TensorStore
from ..storage.store import ShardStore


@dataclass
class Services:
    shard_store: ShardStore
    tensor_store: TensorStore
    # Called as barrier(job_id, func_id) by a worker that published its replica and must wait for
    # the merged model (the HTTP POST /next/{funcId} of network.py:395-414, train/api.go:100-126).
    barrier: Optional[Callable[[str, int], None]] = None
    # Device-resident sync engine (kubeml_b200.parallel.device_sync.DeviceSync) when the job runs
    # on GPUs with peer memory; None selects the host tensor-store path.
    device_sync: Any = None
    # Extra knobs for the worker (fault injection, timers)
    options: dict = field(default_factory=dict)
