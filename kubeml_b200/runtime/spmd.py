"""Process-per-GPU execution under `torch.distributed.run`.

The reference reaches its N function pods through the Fission router
(ml/pkg/train/function.go:103-128).  Here the N workers are the N ranks torchrun started: rank 0
also hosts the train job and posts each fan-out on a key-value bus with worker heartbeats (`StoreBus`);
every rank runs its slice of it on its own GPU and posts its result back.  Ranks beyond the epoch's
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


class WorkerLost(KubeMLException):
    def __init__(self, rank: int, why: str):
        super().__init__(f"worker rank {rank} is not responding ({why})", 503)
        self.rank = rank


class StoreBus:
    """Control channel of the process-per-GPU runtime over the job's key-value store (the TCPStore every
    rank already shares for rendezvous), with worker liveness.

    The reference reaches its function pods through HTTP calls that fail when a pod dies, and Kubernetes
    restarts the pod (readiness probes on every job pod, ml/pkg/ps/job_pod.go:172-186; a 10-try start retry,
    ml/pkg/ps/api.go:192-207).  Blocking collectives have neither property: with a gloo broadcast/gather a
    dead rank leaves rank 0 stuck until the process-group timeout.  So fan-outs travel as keys instead:
    `msg/<seq>` posted by rank 0, `res/<seq>/<rank>` posted by each worker, and every rank refreshes
    `hb/<rank>` from a background thread.  Rank 0 polls for the results and gives up on a rank whose
    heartbeat is older than `KUBEML_HEARTBEAT_TIMEOUT_S` (default 15 s): that function then counts as failed
    (`WorkerLost`, HTTP 503) and the job goes on with the others, as the reference lets an epoch stand when
    one function answered (ml/pkg/train/util.go:144-166).  A lost rank stays lost (torchrun owns the
    process; nothing here can respawn it) and is skipped by later fan-outs at once."""

    def __init__(self, rank: int, world: int, store=None, interval: float = 0.5):
        import pickle
        import threading
        import time
        self._pickle, self._time = pickle, time
        self.rank, self.world = rank, world
        self.store = store if store is not None else self._default_store()
        self.timeout = float(os.environ.get("KUBEML_HEARTBEAT_TIMEOUT_S", "15"))
        self.lost: Dict[int, str] = {}
        self._seq = 0
        self._stop = threading.Event()
        self._beat()
        self._thread = threading.Thread(target=self._loop, args=(interval,), daemon=True, name="kubeml-heartbeat")
        self._thread.start()

    @staticmethod
    def _default_store():
        from torch.distributed import distributed_c10d as c10d
        return dist.PrefixStore("kubeml_bus", c10d._get_default_store())

    def _beat(self) -> None:
        self.store.set(f"hb/{self.rank}", repr(self._time.time()))

    def _loop(self, interval: float) -> None:
        while not self._stop.wait(interval):
            try:
                self._beat()
            except Exception:  # noqa: BLE001 - the store went away with the job
                return

    def age(self, rank: int) -> float:
        try:
            if not self.store.check([f"hb/{rank}"]):
                return float("inf")
            return self._time.time() - float(self.store.get(f"hb/{rank}").decode())
        except Exception:  # noqa: BLE001
            return float("inf")

    # ---- rank 0 ----
    def post(self, msg: Any) -> int:
        self._seq += 1
        self.store.set(f"msg/{self._seq}", self._pickle.dumps(msg))
        return self._seq

    def collect(self, seq: int, ranks: List[int]) -> Dict[int, Any]:
        """Results of `ranks` for fan-out `seq`; a rank that stops beating becomes a `WorkerLost` entry."""
        out: Dict[int, Any] = {}
        pending = [r for r in ranks if r not in self.lost]
        for r in ranks:
            if r in self.lost:
                out[r] = WorkerLost(r, self.lost[r])
        delay = 0.0005
        while pending:
            for r in list(pending):
                key = f"res/{seq}/{r}"
                if self.store.check([key]):
                    out[r] = self._pickle.loads(self.store.get(key))
                    self.store.delete_key(key)
                    pending.remove(r)
                elif self.age(r) > self.timeout:
                    self.lost[r] = f"no heartbeat for {self.timeout:.0f} s"
                    log.error("worker rank %d lost: %s", r, self.lost[r])
                    out[r] = WorkerLost(r, self.lost[r])
                    pending.remove(r)
            if pending:
                self._time.sleep(delay)
                delay = min(delay * 1.5, 0.02)
        try:
            self.store.delete_key(f"msg/{seq}")          # every live rank has read it (it answered)
        except Exception:  # noqa: BLE001
            pass
        return out

    # ---- ranks > 0 ----
    def next_message(self) -> Any:
        """Blocks until rank 0 posts the next fan-out; returns None when rank 0 is gone."""
        self._seq += 1
        key = f"msg/{self._seq}"
        delay = 0.0005
        while not self.store.check([key]):
            if self.age(0) > max(self.timeout, 30.0):
                log.error("rank %d: the control plane (rank 0) stopped beating; leaving", self.rank)
                return None
            self._time.sleep(delay)
            delay = min(delay * 1.5, 0.02)
        return self._pickle.loads(self.store.get(key))

    def reply(self, obj: Any) -> None:
        self.store.set(f"res/{self._seq}/{self.rank}", self._pickle.dumps(obj))

    def close(self) -> None:
        self._stop.set()


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
        self.bus = StoreBus(self.rank, self.world) if self.world > 1 else None
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
        """Worker loop of ranks > 0: run what rank 0 posts until it says stop (or disappears)."""
        while True:
            msg = self.bus.next_message()
            if msg is None or msg.get("stop"):
                return
            try:
                res = self._execute(msg)
            except BaseException as e:  # noqa: BLE001 - the worker already maps exceptions; this is the last resort
                res = ({"error": repr(e), "code": 500}, 500)
            self.bus.reply(res)

    # ---- rank 0 ----------------------------------------------------------------------------
    def run(self, msg: dict) -> List[Any]:
        """One fan-out: post it, run rank 0's own share, collect the others.  A lost worker's entry is the
        error pair ({"error": ..., "code": 503}, 503) a failed function would have returned."""
        with self._lock:
            if self.world == 1:
                return [self._execute(msg)]
            seq = self.bus.post(msg)
            mine = self._execute(msg)
            got = self.bus.collect(seq, list(range(1, self.world)))
            out: List[Any] = [mine]
            for r in range(1, self.world):
                v = got[r]
                out.append(({"error": v.message, "code": 503}, 503) if isinstance(v, WorkerLost) else v)
            return out

    def stop(self) -> None:
        if self.rank == 0 and self.world > 1:
            self.bus.post({"stop": True})

    def invoker(self) -> "SpmdInvoker":
        return SpmdInvoker(self)

    def close(self) -> None:
        if self.bus is not None:
            self.bus.close()
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
