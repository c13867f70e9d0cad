"""Process-per-GPU execution under `torch.distributed.run`.

The reference reaches its N function pods through the Fission router
(ml/pkg/train/function.go:103-128).  Here the N workers are the N ranks torchrun started: rank 0
also hosts the train job and broadcasts each fan-out over a gloo control group; every rank runs its
slice of it on its own GPU and the results are gathered back.  Ranks beyond the epoch's
parallelism receive the framework-internal task `join` so that the peer-memory group stays
collectively consistent (kubeml_b200.parallel.device_sync.SpmdDeviceSyncGroup).
"""
from __future__ import annotations

import datetime
import logging
import os
from typing import Any, Callable, Dict, List, Optional

import torch
import torch.distributed as dist

from ..exceptions import KubeMLException
from ..model.store import TensorStore
from ..storage.store import ShardStore, default_store
from ..train.invoker import Invoker, Result, _payload_error
from .functions import FunctionRegistry
from .services import Services
from .worker import Worker

log = logging.getLogger("kubeml.spmd")


class SpmdRuntime:
    def __init__(self, shard_store: Optional[ShardStore] = None,
                 registry: Optional[FunctionRegistry] = None, options: Optional[dict] = None,
                 use_device_sync: Optional[bool] = None):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", str(self.rank)))
        cuda = torch.cuda.is_available()
        if cuda:
            torch.cuda.set_device(self.local_rank)
        if self.world > 1 and not dist.is_initialized():
            dist.init_process_group(backend="nccl" if cuda else "gloo",
                                    timeout=datetime.timedelta(minutes=10),
                                    **({"device_id": torch.device("cuda", self.local_rank)} if cuda else {}))
        self.ctl = dist.new_group(backend="gloo") if self.world > 1 else None
        self.tensor_store = TensorStore()
        self.shard_store = shard_store or default_store()
        self.registry = registry or FunctionRegistry()
        self.group = None
        if use_device_sync is None:
            use_device_sync = cuda
        if use_device_sync:
            from ..parallel.device_sync import SpmdDeviceSyncGroup
            self.group = SpmdDeviceSyncGroup(self.rank, self.world)
        self.services = Services(shard_store=self.shard_store, tensor_store=self.tensor_store,
                                 barrier=None, device_sync=self.group, options=dict(options or {}))
        self.worker = Worker(self.services, self.registry,
                             gpu_id=self.local_rank if cuda else None, name=f"rank{self.rank}")
        import threading
        self._lock = threading.Lock()       # fan-outs of concurrent jobs are serialised

    # ---- collective helpers --------------------------------------------------------------
    def _bcast(self, obj: Any) -> Any:
        if self.world == 1:
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=0, group=self.ctl)
        return box[0]

    def _gather(self, obj: Any) -> Optional[List[Any]]:
        if self.world == 1:
            return [obj]
        out = [None] * self.world if self.rank == 0 else None
        dist.gather_object(obj, out, dst=0, group=self.ctl)
        return out

    def _execute(self, msg: dict) -> Any:
        if msg.get("release"):
            if self.group is not None:
                self.group.release(msg["release"])
            (self.worker.state.get("step_graphs") or {}).clear()
            return None
        for k, v in (msg.get("env") or {}).items():
            os.environ[k] = str(v)
        for k, v in (msg.get("options") or {}).items():
            self.services.options[k] = v
        q = msg["queries"][self.rank] if self.rank < len(msg["queries"]) else None
        if q is None:
            return None
        return self.worker.invoke(msg["function"], q, msg.get("body"))

    # ---- non-zero ranks ------------------------------------------------------------------
    def serve(self) -> None:
        """Worker loop of ranks > 0: run what rank 0 broadcasts until it says stop."""
        while True:
            msg = self._bcast(None)
            if msg is None or msg.get("stop"):
                return
            self._gather(self._execute(msg))

    # ---- rank 0 ----------------------------------------------------------------------------
    def run(self, msg: dict) -> List[Any]:
        with self._lock:
            self._bcast(msg)
            return self._gather(self._execute(msg)) or []

    def stop(self) -> None:
        if self.rank == 0 and self.world > 1:
            self._bcast({"stop": True})

    def invoker(self) -> "SpmdInvoker":
        return SpmdInvoker(self)

    def close(self) -> None:
        if self.group is not None:
            self.group.close()


class SpmdInvoker(Invoker):
    def __init__(self, rt: SpmdRuntime):
        self.rt = rt
        self.services = rt.services
        self.max_parallelism = rt.world
        self.env: Dict[str, Any] = {}        # environment variables every rank sets before running
        self.options: Dict[str, Any] = {}    # picklable entries merged into services.options

    def _msg(self, function, queries, body=None) -> dict:
        return {"function": function, "queries": queries, "body": body, "env": dict(self.env),
                "options": dict(self.options)}

    def _pad(self, queries: List[Optional[Dict[str, Any]]], template: Dict[str, Any]) -> List[Dict[str, Any]]:
        """Idle ranks get a `join` so collective group maintenance stays in lock-step; `joinFor`
        tells them which task the busy ranks run (only a train fan-out has a membership barrier)."""
        out = list(queries)
        collective = self.rt.group is not None
        for r in range(len(out), self.rt.world):
            out.append(dict(template, task="join", funcId=r, joinFor=template.get("task"))
                       if collective else None)
        return out

    def release_job(self, job_id: str) -> None:
        """Free the job's arenas, peer mappings and captured step graphs on every rank."""
        self.rt.run({"release": job_id})

    def invoke_one(self, func_id, function, query, body=None):
        assert func_id == 0, "single invocations run on rank 0"
        msg = self._msg(function, self._pad([query], query), body)
        res = self.rt.run(msg)
        return res[0]

    def invoke_many(self, function, queries, on_done=None):
        msg = self._msg(function, self._pad(list(queries), queries[0]))
        res = self.rt.run(msg)
        out: List[Result] = []
        for r, q in enumerate(queries):
            payload, status = res[r]
            err = _payload_error(payload, status)
            ok = err is None
            out.append((r, payload if ok else None, err))
            if on_done is not None:
                on_done(r, ok)
        return out
