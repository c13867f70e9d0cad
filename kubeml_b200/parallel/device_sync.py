"""Device-resident K-AVG synchronisation: the GPU replacement for RedisAI + the train job's
merger.

What each verb of the worker loop (python/kubeml/kubeml/network.py:252-310) becomes:

  reference                                   here
  ------------------------------------------  ---------------------------------------------------
  init: TENSORSET every layer  (:444-461)     rank 0 binds the module into its arena
  load_model: TENSORGET + load_state_dict     no-op: the merged model is already in every arena
  reset optimizer state        (:121-128)     momentum buffer zeroed inside the merge kernel
  save_model: .cpu() + TENSORSET per layer    no-op: parameters *are* the replica
  POST /next, block for merge  (:395-414)     csrc/merge.cu launched on the worker's own stream with
                                              a flag barrier in peer memory; no host round-trip
  job: Update x N, Average, Save              the same kernel: peer loads, fp32 sum, /P, peer stores

Which workers take part in round i of an epoch is a pure function of (D, N, K, B)
(`util.round_schedule`), so every rank derives the participant mask locally; that reproduces the
shrinking WaitGroup of ml/pkg/train/job.go:414-438 without the job being involved.

`RankSync` is the per-worker engine.  Two group flavours build the peer tables it needs:
`ThreadedDeviceSyncGroup` (workers are threads of one process; arenas are made visible with
cuMemSetAccess) and `SpmdDeviceSyncGroup` (one process per GPU; arenas travel as POSIX fds).
"""
from __future__ import annotations

import ctypes
import logging
import os
import threading
from typing import Dict, List, Optional

import torch

from .. import _native
from ..exceptions import MergeError, WorkerFault
from ..util import round_schedule
from .arena import ArenaLayout, WeightArena

log = logging.getLogger("kubeml.device_sync")


def _bits(mask: int) -> List[int]:
    return [r for r in range(32) if (mask >> r) & 1]


class PeerTable:
    """Addresses, valid on *this* rank's GPU, of every known rank's arena regions."""

    def __init__(self):
        self.weights: Dict[int, int] = {}
        self.counters: Dict[int, int] = {}
        self.flags: Dict[int, int] = {}
        self.mcast: int = 0
        self.mcast_ranks: frozenset = frozenset()

    def add(self, rank: int, base: int, layout: ArenaLayout) -> None:
        self.weights[rank] = base + layout.off_f32
        self.counters[rank] = base + layout.off_i64
        self.flags[rank] = base + layout.off_flags

    def mask(self) -> int:
        m = 0
        for r in self.weights:
            m |= 1 << r
        return m


class RankSync:
    name = "device"

    def __init__(self, group: "DeviceSyncGroup", rank: int):
        self.group = group
        self.rank = rank
        self.arena: Optional[WeightArena] = None
        self.layout: Optional[ArenaLayout] = None
        self.peers = PeerTable()
        self.merge_seconds = 0.0
        self.merges = 0
        self._events: List[tuple] = []
        self._schedule: List[List[int]] = []
        self._epoch = 0
        self._model = None
        self._write_mask = 0
        self._fused_steps = 0
        self._seq = 0                 # job-wide number of this fan-out: generations never repeat
        self._dead = 0                # ranks this rank has given up on for the current epoch
        self._status_host: Optional[torch.Tensor] = None
        self._round_launched = -1
        self.recovered = 0            # merges re-run without a missing rank

    # ---- init task -----------------------------------------------------------------------
    def publish_reference(self, job_id: str, network: torch.nn.Module) -> List[str]:
        device = next(network.parameters()).device
        if device.type != "cuda":
            device = torch.device("cuda", torch.cuda.current_device())
            network.to(device)
        self._ensure_arena(network, device, copy_in=True)
        torch.cuda.current_stream(device).synchronize()
        self.group.set_reference_holder(self.rank)
        return list(self.layout.order)

    # ---- per task ------------------------------------------------------------------------
    def attach(self, kube_model) -> None:
        """Bind the invocation's fresh module to this rank's arena and agree on membership."""
        self._model = kube_model
        a = kube_model.args
        net = kube_model._network
        joined = self.arena is None
        self._ensure_arena(net, kube_model.device, copy_in=False)
        if joined or (a._task == "train" and getattr(self, "_stale", False)):
            # a worker that joins the job, or one that used its replica as scratch while idle (pre-warm),
            # starts from the merged model
            self.group.copy_reference_into(self)
            self._stale = False
        self.arena.bind(net, copy_in=False)      # also refreshes the bf16 operand copy
        join_for = (kube_model._inv.query or {}).get("joinFor") if a._task == "join" else None
        if a._task == "train" or join_for == "train":
            self._epoch = int(a.epoch)
            q = kube_model._inv.query or {}
            self._seq = int(q.get("syncSeq", self._epoch))
            self._dead = 0
            self._round_launched = -1
            self.arena.status.zero_()
            self._clear_dead_mark()
            self._schedule = (round_schedule(kube_model._dataset.num_docs, a._N, a._K, a.batch_size)
                              if a._task == "train" else [])
            self.merge_seconds = 0.0
            self.merge_wait_seconds = 0.0
            self.arena.flags[48:52].zero_()
            self.epoch_merges = 0
            self.epoch_fused_steps = 0
            self._events = []
            # every rank of this epoch has its arena registered once the barrier opens
            self.group.epoch_barrier(self, a._N, ("train", self._epoch))
            self._write_mask = self.peers.mask()

    def _ensure_arena(self, network: torch.nn.Module, device: torch.device, copy_in: bool) -> None:
        layout = ArenaLayout(network)
        if self.arena is not None:
            if layout.signature() != self.layout.signature():
                raise MergeError(RuntimeError("network structure changed between invocations"))
            if copy_in:
                self.arena.bind(network, copy_in=True)
            return
        self.layout = layout
        self.arena = WeightArena(layout, device, exportable=self.group.exportable)
        if copy_in:
            self.arena.bind(network, copy_in=True)
        torch.cuda.current_stream(self.arena.device).synchronize()   # peers may read right after
        self.group.register(self)

    # ---- per round -----------------------------------------------------------------------
    def load_reference(self, job_id: str, network: torch.nn.Module) -> None:
        opt = getattr(self._model, "optimizer", None)
        if opt is not None and hasattr(opt, "attach_arena") and getattr(opt, "_arena", None) is not self.arena:
            opt.attach_arena(self.arena)

    def can_fuse_last_step(self, optimizer) -> bool:
        """The last local step of a round may be left to the merge kernel's phase 0:
        our flat SGD then skips that launch (`arm_deferral`), eagerly or inside the captured step."""
        return hasattr(optimizer, "arm_deferral") and os.environ.get("KUBEML_FUSE_LAST_STEP", "1") != "0"

    def publish_replica(self, job_id: str, func_id: int, network: torch.nn.Module) -> None:
        """Parameters already are the replica `{job}:{name}/{func_id}`."""

    # ---- generations -------------------------------------------------------------------------------
    # One number per (fan-out, round, attempt), strictly increasing over the life of a job (modulo 2^32;
    # the kernel compares signed differences): the flags in the signal pads only ever grow, so a repeated
    # number would let a barrier pass without waiting.  The fan-out number comes from the train job
    # (`syncSeq`), so every rank - also one that joins later - derives the same value.
    def _gen(self, round_idx: int, attempt: int = 0) -> int:
        return (((self._seq & 0x3FF) << 22) | (((round_idx + 1) & 0x3FFFF) << 4) | (attempt & 0xF)) & 0xFFFFFFFF

    def _tombstone(self) -> int:
        return (((self._seq & 0x3FF) << 22) | 0x3FFFFF) & 0xFFFFFFFF

    def _pads(self, only_self: bool = False):
        p = _native.MergeParams()
        for r, addr in self.peers.flags.items():
            if not only_self or r == self.rank:
                p.flags[r] = addr
        p.my_rank = self.rank
        return p

    def _clear_dead_mark(self) -> None:
        """A rank that (re-)joins an epoch is alive again: clear its mark in every pad it can reach.
        Runs before the epoch's membership barrier, so every peer sees it before the first merge."""
        if len(self.peers.flags) <= 1:
            return
        stream = torch.cuda.current_stream(self.arena.device)
        _native.check(_native.lib().kml_merge_mark_dead(ctypes.byref(self._pads()), self.rank, 0, 0,
                                                        int(stream.cuda_stream)), "kml_merge_mark_dead")
        stream.synchronize()

    def abandon(self, round_idx: int) -> None:
        """This worker's function failed in round `round_idx` of the epoch (before its merge): tell every
        peer not to wait for it and not to count it - the divisor of the reference's merger is the number
        of notifications it received (ml/pkg/train/job.go:383-398) and a failed function does not block
        the others (function.go:180-190)."""
        if self.arena is None or len(self.peers.flags) <= 1:
            return
        try:
            stream = torch.cuda.current_stream(self.arena.device)
            stream.synchronize()                    # merges already launched have run (or timed out)
            first = max(round_idx, self._round_launched + 1)
            _native.check(_native.lib().kml_merge_mark_dead(ctypes.byref(self._pads()), self.rank,
                                                            self._gen(first), self._tombstone(),
                                                            int(stream.cuda_stream)), "kml_merge_mark_dead")
            stream.synchronize()
            log.warning("rank %d left epoch %d at round %d", self.rank, self._epoch, first)
        except Exception as e:  # noqa: BLE001 - already on an error path
            log.error("rank %d could not announce its failure: %s", self.rank, e)

    def _flush_pending_step(self) -> None:
        opt = getattr(self._model, "optimizer", None)
        if opt is not None and getattr(opt, "_deferred", False):
            opt.flush_deferred()          # this rank sat the round out: no merge kernel took the step

    def arrive(self, job_id: str, func_id: int, round_idx: int) -> None:
        self._merge(round_idx)
        self._flush_pending_step()

    def end_epoch(self, job_id: str, func_id: int, rounds_done: int) -> None:
        if rounds_done > 0:
            self._merge(rounds_done - 1)
        self._flush_pending_step()
        stream = torch.cuda.current_stream(self.arena.device)
        stream.synchronize()
        if rounds_done > 0:
            self.check_round(rounds_done - 1)
            stream.synchronize()
        for e0, e1 in self._events:
            self.merge_seconds += e0.elapsed_time(e1) / 1e3
        self._events = []
        # of which: waiting for the slowest peer to finish its local steps (clocked inside the kernel)
        w = self.arena.flags[48:52].cpu().view(torch.int64)
        self.merge_wait_seconds = float(w[0]) / 1e9
        self.arena.flags[48:52].zero_()

    def check_round(self, round_idx: int) -> None:
        """Called after the host has synchronised with the stream that ran round `round_idx`'s merge.
        A merge that timed out on some ranks touched nothing: record those ranks as dead for the rest of
        the epoch and run the merge again without them (up to 3 times); being declared dead by the
        others, or losing a peer half-way through a merge, fails this worker."""
        if self._status_host is None or round_idx != self._round_launched:
            return
        st = int(self._status_host[0])
        attempt = 0
        while st != 0:
            self.arena.status.zero_()
            if st & _native.MERGE_EVICTED:
                raise WorkerFault(f"rank {self.rank} was dropped from the merge by its peers (it arrived too late)")
            if st & _native.MERGE_LATE:
                raise WorkerFault(f"a peer vanished in the middle of a merge (ranks {_bits(st & 0xFF)})")
            missing = st & 0xFF
            attempt += 1
            if attempt > 3 or missing == 0:
                raise WorkerFault(f"merge barrier timed out waiting for rank(s) {_bits(missing)}")
            log.warning("rank %d: rank(s) %s missed round %d of epoch %d; merging without them", self.rank,
                        _bits(missing), round_idx, self._epoch)
            self._dead |= missing
            stream = torch.cuda.current_stream(self.arena.device)
            for r in _bits(missing):          # (the kernel that timed out already wrote the marks; this adds the
                #                                tombstone flags so that no later round waits for the rank)
                _native.check(_native.lib().kml_merge_mark_dead(ctypes.byref(self._pads()), r,
                                                                self._gen(round_idx), self._tombstone(),
                                                                int(stream.cuda_stream)), "kml_merge_mark_dead")
            self._merge(round_idx, attempt=attempt, allow_fuse=False)
            stream.synchronize()
            self.recovered += 1
            st = int(self._status_host[0])

    def drain_losses(self, loss) -> float:
        return float(loss)

    # ---- the merge -----------------------------------------------------------------------
    def _merge(self, round_idx: int, attempt: int = 0, allow_fuse: bool = True) -> None:
        parts = self._schedule[round_idx] if round_idx < len(self._schedule) else [self.rank]
        if self.rank not in parts:
            return
        pmask = 0
        for r in parts:
            pmask |= 1 << r
        pmask &= ~self._dead
        opt = getattr(self._model, "optimizer", None)
        p = _native.MergeParams()
        L = self.layout
        for r, addr in self.peers.weights.items():
            p.replica[r] = addr
            p.ireplica[r] = self.peers.counters[r]
            p.flags[r] = self.peers.flags[r]
        p.n_f32 = L.n_f32
        p.n_i64 = L.n_i64
        p.n_sgd = L.n_param
        p.my_rank = self.rank
        p.participants = pmask
        p.write_mask = (self._write_mask | pmask) & ~self._dead
        single = len(parts) == 1 and (self._write_mask | pmask) == pmask
        p.generation = 0 if single else self._gen(round_idx, attempt)
        use_mc = (self.peers.mcast and frozenset(parts) == self.peers.mcast_ranks and not self._dead
                  and p.write_mask == pmask and os.environ.get("KUBEML_MERGE_NVLS", "1") != "0")
        p.mcast = self.peers.mcast if use_mc else None
        p.status = self.arena.ptr(L.off_status)
        tmo = os.environ.get("KUBEML_MERGE_TIMEOUT_MS")
        tmo_s = float(tmo) / 1e3 if tmo else float(os.environ.get("KUBEML_MERGE_TIMEOUT_S", "20"))
        p.timeout_cycles = int(tmo_s * 1.9e9)
        is_adam = opt is not None and "betas" in getattr(opt, "defaults", {})
        if opt is not None and getattr(opt, "_deferred", False) and allow_fuse:
            g = opt.param_groups[0]
            p.fuse_sgd = 2 if is_adam else 1
            p.grad = self.arena.ptr(L.off_grad)
            p.lr, p.weight_decay = g["lr"], g["weight_decay"]
            if is_adam:
                p.beta1, p.beta2, p.eps = g["betas"][0], g["betas"][1], g["eps"]
            else:
                p.momentum = g["momentum"]
            opt._deferred = False
            p.n_fuse = int(getattr(opt, "_deferred_lo", None) or 0)     # 0: the merge applies the whole step
            opt._deferred_lo = None
            self._fused_steps += 1
            self.epoch_fused_steps = getattr(self, "epoch_fused_steps", 0) + 1
        uses_momentum = opt is not None and bool(getattr(opt, "defaults", {}).get("momentum", 0))
        p.momentum_buf = self.arena.ptr(L.off_state1) if (uses_momentum or is_adam) else None
        if is_adam:                          # both moments and the step count are reset by the merge
            p.state2 = self.arena.ptr(L.off_state2)
            p.step_dev = self.arena.opt_step.data_ptr()
        if opt is not None and hasattr(opt, "_state_clean"):
            opt._state_clean = True           # reset_state() at the start of the next round has nothing left to do
        p.w_bf16 = self.arena.ptr(L.off_bf16) if self.arena.wbf16 is not None else None
        stream = torch.cuda.current_stream(self.arena.device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        _native.check(_native.lib().kml_kavg_merge(ctypes.byref(p), int(stream.cuda_stream)),
                      "kml_kavg_merge")
        e1.record(stream)
        if not single:
            if self._status_host is None:
                self._status_host = torch.zeros(1, dtype=torch.int32).pin_memory()
            self._status_host.copy_(self.arena.status, non_blocking=True)     # read after the round's host sync
        self._round_launched = round_idx
        self._events.append((e0, e1))
        self.merges += 1
        self.epoch_merges = getattr(self, "epoch_merges", 0) + 1

    def close(self) -> None:
        if self.arena is not None:
            self.arena.close()
            self.arena = None


class DeviceSyncGroup:
    exportable = False

    def register(self, rs: RankSync) -> None: ...
    def set_reference_holder(self, rank: int) -> None: ...
    def copy_reference_into(self, rs: RankSync) -> None: ...
    def epoch_barrier(self, rs: RankSync, n: int, key) -> None: ...


class ThreadedDeviceSyncGroup(DeviceSyncGroup):
    """All workers are threads of one process (the in-process Cluster on a multi-GPU host, tests).
    `services.device_sync` is this object; `KubeModel` asks it for the rank engine of its func id."""

    _by_job: Dict[str, "ThreadedDeviceSyncGroup"] = {}
    _cls_lock = threading.Lock()

    def __init__(self, max_ranks: int):
        self.max_ranks = max_ranks
        self.ranks: Dict[int, RankSync] = {}
        self._lock = threading.RLock()
        self._barriers: Dict[tuple, threading.Barrier] = {}
        self._holder = 0

    @classmethod
    def for_job(cls, job_id: str, max_ranks: int) -> "ThreadedDeviceSyncGroup":
        with cls._cls_lock:
            g = cls._by_job.get(job_id)
            if g is None:
                g = cls(max_ranks)
                cls._by_job[job_id] = g
            return g

    @classmethod
    def drop_job(cls, job_id: str) -> None:
        with cls._cls_lock:
            g = cls._by_job.pop(job_id, None)
        if g is not None:
            g.close()

    def rank_sync(self, rank: int) -> RankSync:
        with self._lock:
            rs = self.ranks.get(rank)
            if rs is None:
                rs = RankSync(self, rank)
                self.ranks[rank] = rs
            return rs

    def register(self, rs: RankSync) -> None:
        lib = _native.lib()
        with self._lock:
            ndev = torch.cuda.device_count()
            for d in range(ndev):
                if d != rs.arena.device.index:
                    _native.check(lib.kml_arena_grant_access(rs.arena.handle, d), "kml_arena_grant_access")
            rs.peers.add(rs.rank, rs.arena.base, rs.layout)
            for other in self.ranks.values():
                if other is rs:
                    continue
                if other.arena is None:
                    continue
                other.peers.add(rs.rank, rs.arena.base, rs.layout)
                rs.peers.add(other.rank, other.arena.base, other.layout)

    def set_reference_holder(self, rank: int) -> None:
        self._holder = rank

    def copy_reference_into(self, rs: RankSync) -> None:
        with self._lock:
            src = self.ranks.get(self._holder)
        if src is None or src.arena is None or src is rs:
            return
        torch.cuda.current_stream(src.arena.device).synchronize()
        rs.arena.weights.copy_(src.arena.weights)
        rs.arena.counters.copy_(src.arena.counters)
        torch.cuda.current_stream(rs.arena.device).synchronize()

    def epoch_barrier(self, rs: RankSync, n: int, key) -> None:
        if n <= 1:
            return
        with self._lock:
            b = self._barriers.get(key)
            if b is None:
                b = threading.Barrier(n)
                self._barriers = {key: b}
        try:
            b.wait(timeout=float(os.environ.get("KUBEML_EPOCH_BARRIER_S", "120")))
        except threading.BrokenBarrierError:
            raise WorkerFault("a worker did not reach the epoch barrier")

    # `services.device_sync` verbs are per rank: KubeModel resolves the rank engine here
    def for_func(self, func_id: int) -> RankSync:
        return self.rank_sync(func_id)

    def engine_for(self, job_id: str, func_id: int) -> RankSync:
        return self.rank_sync(func_id)          # this group object already belongs to one job

    def close(self) -> None:
        for rs in list(self.ranks.values()):
            rs.close()
        self.ranks.clear()


class SpmdDeviceSyncGroup(DeviceSyncGroup):
    """One process per GPU (torchrun, or the control plane's process pool).  Each process holds one
    `RankSync`; arenas are exported as POSIX fds and mapped by every peer (csrc/arena.cpp), and,
    where the fabric supports it, bound to one NVSwitch multicast object so the merge can use
    `multimem.ld_reduce` / `multimem.st`.

    Registration is collective: every rank of the world must reach `register()` for the same job
    (idle ranks of a small-N epoch run the function with task `join`)."""

    exportable = True

    def __init__(self, rank: int, world: int, sock_dir: Optional[str] = None, job_tag: int = 0):
        from .peer import FdExchange
        self.rank, self.world = rank, world
        self.fdx = FdExchange(rank, world, sock_dir)
        self._jobs: Dict[str, RankSync] = {}
        self._members: Dict[str, List[int]] = {}
        self.rs = RankSync(self, rank)          # engine of the most recent job (benchmarks use one job at a time)
        self.rs._imported, self.rs._mcast = [], None
        self._tag = 1000 * (job_tag + 1)
        self._holder = 0

    # A job runs on a set of ranks ("members", physical rank numbers in function-id order): all ranks by
    # default, a GPU subset when several jobs share the box.  Inside a job everything is indexed by the
    # LOGICAL rank = function id = position in `members` (slots of the signal pads, replica tables).
    def set_members(self, job_id: Optional[str], members: List[int]) -> None:
        if job_id:
            self._members[job_id] = list(members)

    def _members_of(self, rs: RankSync) -> List[int]:
        return getattr(rs, "members", None) or list(range(self.world))

    def for_func(self, func_id: int) -> RankSync:
        return self.rs

    def engine_for(self, job_id: str, func_id: int) -> RankSync:
        """One engine (arena, peer mappings, multicast object) per job: a control plane may run
        several jobs with different models over the same worker processes."""
        members = self._members.get(job_id) or list(range(self.world))
        if self.rank not in members or members.index(self.rank) != func_id:
            raise MergeError(RuntimeError(f"function {func_id} of job {job_id} scheduled on rank {self.rank} "
                                          f"(members {members})"))
        rs = self._jobs.get(job_id)
        if rs is None:
            rs = RankSync(self, func_id)
            rs._imported, rs._mcast = [], None
            rs.members = members
            import zlib
            rs._tag = (zlib.crc32(job_id.encode()) & 0x3FFF) * 4096     # per-job tag space of the fd exchange
            self._jobs[job_id] = rs
            self.rs = rs
        return rs

    def release(self, job_id: str) -> None:
        rs = self._jobs.pop(job_id, None)
        self._members.pop(job_id, None)
        if rs is not None:
            self._close_rank(rs)

    def _next_tag(self, rs: Optional[RankSync] = None) -> int:
        if rs is not None and hasattr(rs, "_tag"):
            rs._tag += 1
            return rs._tag
        self._tag += 1
        return self._tag

    def register(self, rs: RankSync) -> None:
        lib = _native.lib()
        dev = rs.arena.device.index or 0
        members = self._members_of(rs)
        rs.peers.add(rs.rank, rs.arena.base, rs.layout)
        if len(members) == 1:
            return
        fd = ctypes.c_int(-1)
        _native.check(lib.kml_arena_export_fd(rs.arena.handle, ctypes.byref(fd)), "kml_arena_export_fd")
        got = self.fdx.all_gather_fds([fd.value], self._next_tag(rs), members=members)
        os.close(fd.value)
        for logical, peer in enumerate(members):
            if peer == self.rank:
                continue
            h = ctypes.c_void_p()
            _native.check(lib.kml_arena_import_fd(dev, got[logical][0], rs.arena.nbytes, ctypes.byref(h)),
                          f"kml_arena_import_fd(rank {peer})")
            os.close(got[logical][0])
            rs._imported.append(h)
            rs.peers.add(logical, int(lib.kml_arena_ptr(h)), rs.layout)
        self._setup_multicast(rs, dev)

    def _setup_multicast(self, rs: RankSync, dev: int) -> None:
        members = self._members_of(rs)
        root = members[0]
        if os.environ.get("KUBEML_MERGE_NVLS", "1") == "0":
            self.fdx.barrier(self._next_tag(rs), members=members)
            return
        lib = _native.lib()
        sup = ctypes.c_int(0)
        ok = lib.kml_mcast_supported(dev, ctypes.byref(sup)) == 0 and sup.value != 0
        m = ctypes.c_void_p()
        fds: List[int] = []
        if self.rank == root and ok:
            if lib.kml_mcast_create(len(members), rs.arena.nbytes, ctypes.byref(m)) == 0:
                fd = ctypes.c_int(-1)
                if lib.kml_mcast_export_fd(m, ctypes.byref(fd)) == 0:
                    fds = [fd.value]
            else:
                log.warning("multicast create failed: %s", lib.kml_last_error().decode())
        got = self.fdx.broadcast_fds(fds, root, self._next_tag(rs), members=members)
        for f in fds:
            os.close(f)
        have = bool(got) and ok
        if have and self.rank != root:
            if lib.kml_mcast_import_fd(got[0], len(members), rs.arena.nbytes, ctypes.byref(m)) != 0:
                have = False
            os.close(got[0])
        if have and lib.kml_mcast_add_device(m, dev) != 0:
            log.warning("multicast add_device failed: %s", lib.kml_last_error().decode())
            have = False
        # everyone must have added its device before anyone binds
        self.fdx.barrier(self._next_tag(rs), members=members)
        mc_ptr = ctypes.c_void_p()
        if have and lib.kml_mcast_bind(m, dev, rs.arena.handle, ctypes.byref(mc_ptr)) != 0:
            log.warning("multicast bind failed: %s", lib.kml_last_error().decode())
            have = False
        have = self.fdx.all_agree(have, self._next_tag(rs), members=members)
        if have:
            rs._mcast = m
            rs.peers.mcast = int(mc_ptr.value) + rs.layout.off_f32
            rs.peers.mcast_ranks = frozenset(range(len(members)))
            log.info("rank %d: NVLS multicast mapping ready (%d GPUs)", self.rank, len(members))

    def set_reference_holder(self, rank: int) -> None:
        self._holder = rank

    def copy_reference_into(self, rs: RankSync) -> None:
        if len(self._members_of(rs)) == 1 or rs.rank == self._holder:
            return
        src = rs.peers.weights.get(self._holder)
        if src is None:
            return
        from .arena import _DeviceSpan
        L = rs.layout
        w = torch.as_tensor(_DeviceSpan(src, (L.n_f32,), "<f4", self), device=rs.arena.device)
        rs.arena.weights.copy_(w)
        if L.n_i64:
            c = torch.as_tensor(_DeviceSpan(rs.peers.counters[self._holder], (L.n_i64,), "<i8", self),
                                device=rs.arena.device)
            rs.arena.counters[: L.n_i64].copy_(c)
        torch.cuda.current_stream(rs.arena.device).synchronize()

    def epoch_barrier(self, rs: RankSync, n: int, key) -> None:
        members = self._members_of(rs)
        if len(members) > 1:
            self.fdx.barrier(self._next_tag(rs), members=members)

    def _close_rank(self, rs: RankSync) -> None:
        lib = _native.lib()
        if rs.arena is not None:
            torch.cuda.synchronize(rs.arena.device)
        if getattr(rs, "_mcast", None) is not None:
            lib.kml_mcast_destroy(rs._mcast)
            rs._mcast = None
        for h in getattr(rs, "_imported", []):
            lib.kml_arena_destroy(h)
        rs._imported = []
        rs.close()

    def close(self) -> None:
        for jid in list(self._jobs):
            self.release(jid)
        self._close_rank(self.rs)
        self.fdx.close()
