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
import threading
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
    `mb/<rank>/<k>` (a mailbox per rank) posted by rank 0, `res/<rank>/<k>` posted by the worker, and every rank refreshes
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
        self._sent: Dict[int, int] = {}
        self._post_lock = threading.Lock()
        self._stop = threading.Event()
        self._t0 = time.time()
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
                # never seen: the process may still be starting (importing torch takes a while) - count from
                # this bus's own start, with a minute's grace
                return max(0.0, self._time.time() - self._t0 - 60.0)
            return self._time.time() - float(self.store.get(f"hb/{rank}").decode())
        except Exception:  # noqa: BLE001
            return float("inf")

    # ---- rank 0 ----
    def post(self, msg: Any, ranks: List[int]) -> Dict[int, int]:
        """Puts `msg` into the mailbox of every rank in `ranks`; returns the ticket (mailbox position) per
        rank.  Mailboxes are per rank so that fan-outs of different jobs to disjoint ranks do not order
        against each other."""
        blob = self._pickle.dumps(msg)
        tickets: Dict[int, int] = {}
        with self._post_lock:
            for r in ranks:
                self._sent[r] = self._sent.get(r, 0) + 1
                tickets[r] = self._sent[r]
        for r, k in tickets.items():
            if r not in self.lost:
                self.store.set(f"mb/{r}/{k}", blob)
        return tickets

    def collect(self, tickets: Dict[int, int]) -> Dict[int, Any]:
        """Results for the given tickets; a rank that stops beating becomes a `WorkerLost` entry."""
        out: Dict[int, Any] = {}
        pending = [r for r in tickets if r not in self.lost]
        for r in tickets:
            if r in self.lost:
                out[r] = WorkerLost(r, self.lost[r])
        delay = 0.0005
        while pending:
            for r in list(pending):
                key = f"res/{r}/{tickets[r]}"
                if self.store.check([key]):
                    out[r] = self._pickle.loads(self.store.get(key))
                    for k in (key, f"mb/{r}/{tickets[r]}"):
                        try:
                            self.store.delete_key(k)
                        except Exception:  # noqa: BLE001
                            pass
                    pending.remove(r)
                elif self.age(r) > self.timeout:
                    self.lost[r] = f"no heartbeat for {self.timeout:.0f} s"
                    log.error("worker rank %d lost: %s", r, self.lost[r])
                    out[r] = WorkerLost(r, self.lost[r])
                    pending.remove(r)
            if pending:
                self._time.sleep(delay)
                delay = min(delay * 1.5, 0.02)
        return out

    # ---- ranks > 0 ----
    def next_message(self) -> Any:
        """Blocks until rank 0 posts into this rank's mailbox; returns None when rank 0 is gone."""
        self._seq += 1
        key = f"mb/{self.rank}/{self._seq}"
        delay = 0.0005
        while not self.store.check([key]):
            if self.age(0) > max(self.timeout, 30.0):
                log.error("rank %d: the control plane (rank 0) stopped beating; leaving", self.rank)
                return None
            self._time.sleep(delay)
            delay = min(delay * 1.5, 0.02)
        return self._pickle.loads(self.store.get(key))

    def reply(self, obj: Any) -> None:
        self.store.set(f"res/{self.rank}/{self._seq}", self._pickle.dumps(obj))

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
        if self.group is None:
            # no GPUs: the merge is an all-reduce over the control group (static parallelism == world size only)
            from ..train.sync import CollectiveSync
            self.services.options.setdefault("sync_factory", lambda: CollectiveSync(self.tensor_store))
        self.worker = Worker(self.services, self.registry,
                             gpu_id=self.local_rank if cuda else None, name=f"rank{self.rank}")
        self._lock = threading.Lock()       # rank 0's own share of a fan-out
        self._alloc_lock = threading.Lock()
        self._held: set = set()

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
        members = msg.get("members") or list(range(self.world))
        if self.rank not in members:
            return None
        idx = members.index(self.rank)            # the function id this rank runs: its position in the job's GPU set
        q = msg["queries"][idx] if idx < len(msg["queries"]) else None
        if q is None:
            return None
        if self.group is not None:
            self.group.set_members(q.get("jobId"), members)
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
    def run(self, msg: dict, ranks: Optional[List[int]] = None) -> List[Any]:
        """One fan-out to `ranks` (default: every rank): post it, run rank 0's own share if it has one, collect
        the others; results come back in the order of `ranks`.  A lost worker's entry is the error pair
        ({"error": ..., "code": 503}, 503) a failed function would have returned.  Fan-outs to disjoint rank sets
        may run concurrently (one job per GPU subset); two fan-outs that share a rank are serialised by that
        rank's mailbox."""
        ranks = list(range(self.world)) if ranks is None else list(ranks)
        msg = dict(msg, members=msg.get("members") or ranks)
        if self.world == 1:
            return [self._execute(msg)]
        remote = [r for r in ranks if r != 0]
        tickets = self.bus.post(msg, remote)
        mine = None
        if 0 in ranks:
            with self._lock:                      # rank 0's own GPU runs one function at a time
                mine = self._execute(msg)
        got = self.bus.collect(tickets)
        out: List[Any] = []
        for r in ranks:
            v = mine if r == 0 else got[r]
            out.append(({"error": v.message, "code": 503}, 503) if isinstance(v, WorkerLost) else v)
        return out

    # ---- GPU subsets for concurrent jobs (the reference runs one pod per job, several at once:
    #      ml/pkg/ps/api.go:139-222, ml/pkg/ps/job_pod.go:96-217) --------------------------------------
    def allocate(self, n: int) -> List[int]:
        """Reserves `n` ranks (GPUs) that no other job holds; raises when fewer are free."""
        with self._alloc_lock:
            free = [r for r in range(self.world) if r not in self._held and (self.bus is None or r not in self.bus.lost)]
            if len(free) < n:
                raise KubeMLException(f"{n} GPUs asked for, {len(free)} free", 503)
            take = free[:n]
            self._held.update(take)
            return take

    def free(self, ranks: List[int]) -> None:
        with self._alloc_lock:
            self._held.difference_update(ranks)

    def stop(self) -> None:
        if self.rank == 0 and self.world > 1:
            self.bus.post({"stop": True}, list(range(1, self.world)))

    def invoker(self, ranks: Optional[List[int]] = None) -> "SpmdInvoker":
        """Invoker over all ranks, or over a GPU subset obtained from `allocate()` (function id i of the job
        runs on `ranks[i]`)."""
        return SpmdInvoker(self, ranks)

    def close(self) -> None:
        if self.bus is not None:
            self.bus.close()
        if self.group is not None:
            self.group.close()


class SpmdInvoker(Invoker):
    def __init__(self, rt: SpmdRuntime, ranks: Optional[List[int]] = None):
        self.rt = rt
        self.ranks = list(range(rt.world)) if ranks is None else list(ranks)
        self.services = rt.services
        self.max_parallelism = len(self.ranks)
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
        for r in range(len(out), len(self.ranks)):
            out.append(dict(template, task="join", funcId=r, joinFor=template.get("task"))
                       if collective else None)
        return out

    def release_job(self, job_id: str) -> None:
        """Free the job's arenas, peer mappings and captured step graphs on every rank."""
        self.rt.run({"release": job_id}, self.ranks)

    def invoke_one(self, func_id, function, query, body=None):
        assert func_id == 0, "single invocations run on the job's first rank"
        msg = self._msg(function, self._pad([query], query), body)
        res = self.rt.run(msg, self.ranks)
        return res[0]

    def invoke_many(self, function, queries, on_done=None):
        msg = self._msg(function, self._pad(list(queries), queries[0]))
        res = self.rt.run(msg, self.ranks)
        out: List[Result] = []
        for r, q in enumerate(queries):
            payload, status = res[r]
            err = _payload_error(payload, status)
            ok = err is None
            out.append((r, payload if ok else None, err))
            if on_done is not None:
                on_done(r, ok)
        return out
