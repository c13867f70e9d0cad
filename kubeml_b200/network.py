"""`KubeModel`: the user-facing training abstraction (reference:
python/kubeml/kubeml/network.py:29-476).

Public surface kept: `KubeModel(network, dataset, gpu=False)`; overridables
`configure_optimizers() -> Optimizer`, `init()`, `train(batch, batch_index) -> float`,
`validate(batch, batch_index) -> (accuracy, loss)`, `infer(data)`; proxies `__call__`,
`parameters()`, `apply()`; attributes `lr, batch_size, epoch, optimizer, logger, device, args,
task, platform`; and `start()`, which dispatches on the invocation's `task`
(init | train | val | infer) and returns the function result.

What changed underneath: the invocation arrives through the worker runtime rather
than Flask; weights move through a `sync` engine (host tensor store or the device arena) rather
than RedisAI; `train`/`validate` may return 0-dim tensors, which are accumulated on the device
and read back once per round instead of one `.item()` sync per batch.
"""
from __future__ import annotations

import logging
import os
from abc import ABC
from collections import defaultdict
from typing import Any, Callable, Iterable, List, Sequence, Tuple, Union

import numpy as np
import torch
import torch.nn as nn

from .args import _KubeArgs
from .dataset import KubeDataset
from .exceptions import (DataError, InvalidFormatError, KubeMLException, StorageError)
from .runtime.context import current_invocation
from .util import get_gpu, get_subset_period, split_minibatches

Number = Union[float, torch.Tensor]


class KubeModel(ABC):
    # A subclass whose train() neither synchronises with the host (.item(), prints of tensor
    # values) nor changes shape-dependent control flow between batches may set this: the worker then
    # captures one training step in a CUDA graph and replays it (launch-bound small models).
    graph_safe = False

    def __init__(self, network: nn.Module, dataset: KubeDataset, gpu: bool = False):
        self._network = network
        self._dataset = dataset
        self.platform = "gpu" if gpu else "cpu"
        self.device = None
        self.args = None
        self.logger = None

        # refreshed from the invocation arguments on every call (network.py:43-49)
        self.lr = None
        self.batch_size = None
        self.task = None
        self.optimizer = None
        self.epoch = None

        self._sync = None
        self._inv = None
        # device-timed seconds spent in merges during the last train task (history extension)
        self.merge_seconds = 0.0
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.train_seconds = 0.0      # device-timed: first batch staged -> last merge done
        self.steps = 0
        self.gpu_launches = 0

    # ---- proxies to the wrapped module (network.py:55-89) -------------------------------------
    def __call__(self, *args, **kwargs):
        return self._network(*args, **kwargs)

    def parameters(self):
        return self._network.parameters()

    def apply(self, fn: Callable[[nn.Module], None]):
        self._network.apply(fn)
        return self

    # ---- invocation plumbing ---------------------------------------------------------------
    def _read_args(self) -> None:
        self.args = _KubeArgs.parse()
        self.lr = self.args.lr
        self.batch_size = self.args.batch_size
        self.task = self.args._task
        self.epoch = self.args.epoch

    def _get_logger(self) -> None:
        self.logger = self._inv.logger

    def _make_sync(self):
        sv = self._inv.services
        if sv is None:
            raise KubeMLException("invocation carries no services", 500)
        if sv.device_sync is not None and self.platform == "gpu" and torch.cuda.is_available():
            ds = sv.device_sync
            if hasattr(ds, "engine_for"):
                return ds.engine_for(self.args._job_id, self.args._func_id)
            return ds
        factory = (sv.options or {}).get("sync_factory")
        if factory is not None:          # e.g. CollectiveSync: K-AVG through torch.distributed
            ws = self._inv.worker_state if self._inv.worker_state is not None else {}
            if "sync_engine" not in ws:
                ws["sync_engine"] = factory()
            return ws["sync_engine"]
        from .train.sync import HostSync
        return HostSync(sv.tensor_store, sv.barrier)

    def start(self) -> Tuple[Any, int]:
        """Run the task named by the invocation; returns `(json_payload, 200)` like the
        reference's `(flask.Response, 200)` (network.py:146-172)."""
        self._inv = current_invocation()
        self._read_args()
        self._get_logger()
        self._sync = self._make_sync()

        if self.task == "init":
            return self.__initialize(), 200
        if self.task == "train":
            try:
                loss = self.__train()
            except BaseException:
                # a failed function must not hold up the others: tell the peers this rank is out for the
                # rest of the epoch (the reference's deferred finish path, ml/pkg/train/function.go:180-190)
                abandon = getattr(self._sync, "abandon", None)
                if abandon is not None:
                    abandon(getattr(self, "_round_idx", 0))
                raise
            return {"loss": loss, "merge_seconds": self.merge_seconds,
                    "h2d_bytes": self.h2d_bytes, "d2h_bytes": self.d2h_bytes,
                    "train_seconds": self.train_seconds, "steps": self.steps,
                    "gpu_launches": self.gpu_launches, "merges": int(getattr(self, "merges", 0)),
                    "merge_wait_seconds": float(getattr(self._sync, "merge_wait_seconds", 0.0)),
                    "fused_last_steps": int(getattr(self, "fused_last_steps", 0)),
                    **({"trace": self.trace} if getattr(self, "trace", None) else {})}, 200
        if self.task == "val":
            acc, loss, length = self.__validate()
            return {"loss": loss, "accuracy": acc, "length": length}, 200
        if self.task == "infer":
            return {"predictions": self.__infer()}, 200
        if self.task == "checkpoint":
            # framework-internal: persist the device-resident reference model
            self._set_device()
            self._sync.attach(self)
            from .utils.checkpoint import save_state_dict
            path = save_state_dict(self._inv.query["ckptDir"], self.args._job_id, self._network.state_dict())
            return {"path": path}, 200
        if self.task == "join":
            # framework-internal: an idle rank of a process-per-GPU group keeps its arena mapped
            # into the peers and takes part in the epoch's membership barrier
            self._set_device()
            self._sync.attach(self)
            if (self._inv.query or {}).get("joinFor") == "train":
                self._prewarm_join()
            return {}, 200
        raise KubeMLException(f"Task {self.task} not recognized", 400)

    def _prewarm_join(self) -> None:
        """An idle worker of a growing job gets ready while the active ones train: the first time it is told to
        `join` a training epoch it runs the user's train() on one batch - the eager warm-up calls and the CUDA
        graph captures (with and without the deferred optimizer step) that its first real epoch would otherwise
        pay while its peers wait for it in the merge.  Its replica is scratch until then: the sync engine copies
        the merged model again when the worker becomes active (`RankSync._stale`)."""
        ws = self._inv.worker_state if self._inv.worker_state is not None else {}
        done = ws.setdefault("prewarmed", set())
        job = self.args._job_id
        if (job in done or self.device is None or self.device.type != "cuda" or not self._graphs_enabled()
                or os.environ.get("KUBEML_PREWARM_JOIN", "1") == "0" or self._dataset.num_docs <= 0):
            return
        done.add(job)
        try:
            self._network.train()
            self._config_optimizer()
            self._sync.load_reference(job, self._network)
            ds = self._dataset
            ds._want_pinned, ds._device_resident, ds._trace = False, False, False
            ds._load_train_data(start=0, end=min(self._dataset.num_docs, -(-self.batch_size // 64)))
            batches = ds._batches(self.batch_size, self.device)
            batch = next(iter(batches))
            can_fuse = getattr(self._sync, "can_fuse_last_step", None)
            variants = [False] + ([True] if can_fuse is not None and can_fuse(self.optimizer) else [])
            for defer in variants:
                for _ in range(4):            # eager call(s), the capture, replay(s)
                    self._graphed_step(batch, 0, defer and self._single_step is True)
                    if getattr(self.optimizer, "_deferred", False):
                        self.optimizer._deferred = False
            torch.cuda.synchronize(self.device)
            self._sync._stale = True
            self.logger.info("function %d pre-warmed for job %s", self.args._func_id, job)
        except Exception as e:  # noqa: BLE001 - a failed warm-up only costs time later
            self.logger.warning("pre-warm of function %d failed: %s", self.args._func_id, e)
            self._sync._stale = True

    # ---- optimizer (network.py:99-138) ---------------------------------------------------------
    def _config_optimizer(self) -> None:
        self.optimizer = self.configure_optimizers()
        # A native network keeps bf16 tensor-core copies of its weights that only the flat (arena)
        # optimizers maintain: on the host-sync path give it its private arena up front.
        net = self._network
        if getattr(net, "native", False) and hasattr(self.optimizer, "attach_arena") \
                and getattr(self._sync, "arena", None) is None:
            net._ensure_arena()
            self.optimizer.attach_arena(net._arena)

    def _reset_optimizer_state(self) -> None:
        """Momentum & co. are discarded at every sync: state carried over from a pre-merge
        model "can hinder progress" (network.py:121-128)."""
        if self.optimizer is None:
            return
        reset = getattr(self.optimizer, "reset_state", None)
        if callable(reset):
            reset()
        else:
            self.optimizer.state = defaultdict(dict)

    # ---- task: init ------------------------------------------------------------------------
    def __initialize(self) -> List[str]:
        # The library's hook is `init(self)` (network.py:466); the reference's README example defines
        # `init(self, model)` (README.md:191): accept both.
        import inspect
        try:
            takes_model = len(inspect.signature(self.init).parameters) >= 1
        except (TypeError, ValueError):
            takes_model = False
        if takes_model:
            self.init(self._network)
        else:
            self.init()
        q = self._inv.query or {}
        if q.get("resumeFrom"):
            # warm start: the checkpoint of an earlier task replaces the function's initialisation before the
            # reference model is published (to the tensor store or, on GPUs, into rank 0's arena)
            from .utils.checkpoint import load_checkpoint
            state, _ = load_checkpoint(q["ckptDir"], q["resumeFrom"])
            missing = set(self._network.state_dict()) ^ set(state)
            if missing:
                raise StorageError(KeyError(f"checkpoint {q['resumeFrom']} does not match the network: {sorted(missing)[:4]}"))
            self._network.load_state_dict(state)
        try:
            return self._sync.publish_reference(self.args._job_id, self._network)
        except KeyError as e:
            raise StorageError(e)

    # ---- hooks (network.py:191-224) ------------------------------------------------------------
    def _on_train_start(self) -> None:
        self._set_device()
        self._network.train()
        self._sync.attach(self)
        self._config_optimizer()

    def _on_train_end(self) -> None:
        pass

    def _on_iteration_start(self) -> None:
        self._sync.load_reference(self.args._job_id, self._network)
        self._reset_optimizer_state()

    def _on_iteration_end(self) -> None:
        self._sync.publish_replica(self.args._job_id, self.args._func_id, self._network)

    def _batch_to_device(self, batch):
        """network.py:226-250: tensors, tuples and sequences of tensors are moved; anything else
        is passed through."""
        if isinstance(batch, torch.Tensor):
            return batch.to(self.device, non_blocking=True)
        if isinstance(batch, tuple):
            return tuple(_to(e, self.device) for e in batch)
        if isinstance(batch, Sequence) and not isinstance(batch, (str, bytes)):
            return type(batch)([_to(e, self.device) for e in batch])
        return batch

    # ---- task: train (network.py:252-310) ------------------------------------------------------
    def __train(self) -> float:
        self._on_train_start()
        a = self.args
        assigned = split_minibatches(range(self._dataset.num_docs), a._N)[a._func_id]
        per_iter = get_subset_period(a._K, a.batch_size, assigned)
        self.logger.debug("Subsets per iteration: %s", per_iter)
        intervals = range(assigned.start, assigned.stop, per_iter) if per_iter > 0 else range(0)

        opts = self._inv.services.options or {}
        fault = _fault_hook(opts.get("inject_fault") or os.environ.get("KUBEML_INJECT_FAULT"))
        on_cuda = self.device.type == "cuda"
        loss_host = 0.0
        loss_dev: Number = 0.0
        num_iterations = 0
        self.merge_seconds = 0.0
        self.h2d_bytes = self.d2h_bytes = 0
        self._dataset._want_pinned = on_cuda and self._dataset._has_batch_transform()
        self._dataset._device_resident = bool(opts.get("device_resident_data"))
        tracing = on_cuda and bool(opts.get("trace") or os.environ.get("KUBEML_TRACE") == "1")
        self._dataset._trace = tracing
        trace_steps, trace_h2d = [], []
        # per-step results are copied device->host asynchronously into page-locked memory and read
        # after the round's merge was queued: one sync per round, not one `.item()` per batch
        stage = _LossStage(self.device) if on_cuda else None
        launches0 = _native_launches()
        graph_l0 = (self._inv.worker_state or {}).get("graph_launches", 0)
        t_ev0 = t_ev1 = None
        if on_cuda:
            torch.cuda.synchronize(self.device)
            t_ev0, t_ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t_ev0.record()
        import time as _time
        t_wall = _time.perf_counter()
        self._round_idx = 0

        def _prepare(start):
            self._dataset._load_train_data(start=start, end=min(assigned.stop, start + per_iter))
            return self._dataset._batches(self.batch_size, self.device)

        # The host reads the round's losses and the merge status once per round, in step with the stream.
        # Everything the next round needs from the host (shard lookup, batch views) is prepared *before* that
        # wait, while the GPU is still busy with this round, so that the only thing between the end of the
        # merge and the next round's first kernel is one graph launch.
        prefetch = on_cuda and os.environ.get("KUBEML_ROUND_PREFETCH", "1") != "0"
        # ... and that launch is already queued: the host hands the next round's first step to the stream
        # *before* it waits for this round's merge, and reads the losses and the merge status while that step
        # runs.  If the status then says the merge timed out (a peer went silent - the rare path), the repeat
        # merge of `check_round` runs behind that one extra local step.
        lookahead = prefetch and os.environ.get("KUBEML_ROUND_LOOKAHEAD", "1") != "0"
        # KUBEML_NVTX=1: ranges for an external timeline tool (Nsight Systems) around the epoch, every round and
        # every merge launch; the built-in timeline is `options["trace"]` (device spans in the job's Chrome trace)
        nvtx = torch.cuda.nvtx if (on_cuda and os.environ.get("KUBEML_NVTX") == "1") else None
        if nvtx is not None:
            nvtx.range_push(f"kubeml epoch {a.epoch} job {a._job_id} func {a._func_id}")
        starts = list(intervals)
        ready = None
        pending = None          # (round, event after its merge, losses pushed up to it) not read back yet

        def settle() -> float:
            nonlocal pending
            r, ev, n = pending
            pending = None
            v = stage.drain(n, ev)
            check = getattr(self._sync, "check_round", None)
            if check is not None:
                check(r)
            return v
        for round_idx, i in enumerate(starts):
            self._round_idx = round_idx
            if nvtx is not None:
                nvtx.range_push(f"round {round_idx}")
            batches = ready if ready is not None else _prepare(i)
            ready = None
            num_iterations += len(batches)

            self._on_iteration_start()
            graphing = self._graphs_enabled()
            can_fuse = getattr(self._sync, "can_fuse_last_step", None)
            fuse_last = bool(can_fuse is not None and can_fuse(self.optimizer))
            sliced = isinstance(batches, _sliced_type())
            for idx, batch in enumerate(batches):
                if pending is not None and idx == 1:
                    loss_host += settle()
                if fault and fault(a._func_id, a.epoch, round_idx, idx):
                    raise KubeMLException(f"injected fault in function {a._func_id}", 500)
                if not sliced:
                    batch = self._batch_to_device(batch)
                # the round's last optimizer step is left to the merge kernel - but only for functions
                # that were seen to call step() exactly once per batch
                defer = fuse_last and idx == len(batches) - 1 and self._single_step is True
                if tracing:
                    s0 = torch.cuda.Event(enable_timing=True)
                    s0.record()
                step_loss = self._graphed_step(batch, idx, defer) if graphing else self._eager_step(batch, idx, defer)
                if tracing:
                    s1 = torch.cuda.Event(enable_timing=True)
                    s1.record()
                    trace_steps.append((round_idx, idx, s0, s1))
                if stage is not None and isinstance(step_loss, torch.Tensor) and step_loss.is_cuda:
                    stage.push(step_loss)
                elif isinstance(step_loss, torch.Tensor):
                    loss_dev = loss_dev + step_loss
                else:
                    loss_host += float(step_loss)
            if pending is not None:
                loss_host += settle()
            self._on_iteration_end()
            self.h2d_bytes += getattr(batches, "h2d_bytes", 0)
            if tracing and getattr(batches, "trace", None):
                trace_h2d += [(round_idx, b, e0, e1) for b, e0, e1 in batches.trace]

            # every round but the last blocks on the merge barrier (network.py:305-306)
            if nvtx is not None:
                nvtx.range_pop()
            if i != intervals[-1]:
                if nvtx is not None:
                    nvtx.range_push(f"merge {round_idx}")
                self._sync.arrive(a._job_id, a._func_id, round_idx)
                if nvtx is not None:
                    nvtx.range_pop()
            if prefetch and round_idx + 1 < len(starts):
                ready = _prepare(starts[round_idx + 1])
                if hasattr(ready, "prestage"):
                    ready.prestage()          # first batch on its way to the device before the wait below
            if stage is not None:
                if lookahead and ready is not None:
                    ev = torch.cuda.Event()
                    ev.record()
                    pending = (round_idx, ev, stage.n)
                else:
                    loss_host += stage.drain()
                    check = getattr(self._sync, "check_round", None)
                    if check is not None:        # the host is in step with the stream here: did the merge go through?
                        check(round_idx)

        merge_events = list(getattr(self._sync, "_events", [])) if tracing else []
        self._sync.end_epoch(a._job_id, a._func_id, len(intervals))
        if nvtx is not None:
            nvtx.range_pop()
        if on_cuda:
            t_ev1.record()
            torch.cuda.synchronize(self.device)
            self.train_seconds = t_ev0.elapsed_time(t_ev1) / 1e3
            if tracing:
                # device timeline of this worker's epoch, milliseconds from the epoch's first event: compute
                # steps, host->device copies (copy stream) and merges - what shows the copy of batch i+1 under
                # the compute of batch i
                rel = lambda e0, e1: (round(t_ev0.elapsed_time(e0), 4), round(e0.elapsed_time(e1), 4))
                self.trace = {"step": [(r, b) + rel(e0, e1) for r, b, e0, e1 in trace_steps],
                              "h2d": [(r, b) + rel(e0, e1) for r, b, e0, e1 in trace_h2d],
                              "merge": [(i, 0) + rel(e0, e1) for i, (e0, e1) in enumerate(merge_events)]}
        else:
            self.train_seconds = _time.perf_counter() - t_wall
        self.steps = num_iterations
        ws = self._inv.worker_state or {}
        self.gpu_launches = _native_launches() - launches0 + ws.get("graph_launches", 0) - graph_l0
        if stage is not None:
            self.d2h_bytes = stage.bytes
        self.merge_seconds = float(getattr(self._sync, "merge_seconds", 0.0))
        self.merges = int(getattr(self._sync, "epoch_merges", 0))
        self.fused_last_steps = int(getattr(self._sync, "epoch_fused_steps", 0))
        self._on_train_end()
        if num_iterations == 0:
            return 0.0
        return (loss_host + float(self._sync.drain_losses(loss_dev))) / num_iterations

    # ---- CUDA-graph replay of the user's train() ---------------------------------------------------
    def _graphs_enabled(self) -> bool:
        if self.device is None or self.device.type != "cuda" or not self.graph_safe:
            return False
        if (self._inv.services.options or {}).get("cuda_graphs") is False:
            return False
        return os.environ.get("KUBEML_CUDA_GRAPH", "1") != "0"

    # Does the user's train() call optimizer.step() exactly once per batch?  Observed on eager calls and
    # remembered in the worker's state: later invocations (epochs) only replay the captured step.
    @property
    def _single_step(self):
        ws = self._inv.worker_state if self._inv is not None and self._inv.worker_state is not None else self.__dict__
        return ws.setdefault("single_step", {}).get((self.args._job_id, type(self).__name__))

    @_single_step.setter
    def _single_step(self, v) -> None:
        ws = self._inv.worker_state if self._inv is not None and self._inv.worker_state is not None else self.__dict__
        ws.setdefault("single_step", {})[(self.args._job_id, type(self).__name__)] = v

    def _eager_step(self, batch, idx, defer: bool = False):
        opt = self.optimizer
        c0 = getattr(opt, "_step_calls", None)
        if defer and not opt.arm_deferral():
            defer = False
        try:
            out = self.train(batch, idx)
        finally:
            if defer:
                opt.disarm()
        if c0 is not None:
            once = (opt._step_calls - c0) == 1
            self._single_step = once if self._single_step is None else (self._single_step and once)
        return out

    def _graphed_step(self, batch, idx, defer: bool = False):
        """First two calls per (shapes, hyper-parameters) run eagerly (allocations, lazy inits), the
        third is captured, later ones replay (one eager call once the job has a captured graph already).  The captured program only touches arena memory and
        buffers owned by the capturing KubeModel, which the cache keeps alive, so it stays valid for
        later invocations (epochs) whose fresh module objects alias the same arena."""
        ws = self._inv.worker_state if self._inv.worker_state is not None else {}
        cache = ws.setdefault("step_graphs", {})
        tensors = [t for t in (batch if isinstance(batch, (tuple, list)) else (batch,))
                   if isinstance(t, torch.Tensor)]
        opt = self.optimizer
        hyper = tuple(sorted((k, v) for k, v in getattr(opt, "defaults", {}).items()
                             if isinstance(v, (int, float, bool))))
        lr_now = tuple(g.get("lr") for g in getattr(opt, "param_groups", []))
        key = (self.args._job_id, type(self).__name__, tuple((tuple(t.shape), t.dtype) for t in tensors),
               hyper, lr_now, id(getattr(self._sync, "arena", None)), bool(defer))
        ent = cache.get(key)
        if ent is None:
            ent = {"calls": 0, "graph": None, "failed": False}
            cache[key] = ent
        if ent["graph"] is not None:
            for s_t, t in zip(ent["static"], tensors):
                if s_t.data_ptr() != t.data_ptr():     # a dataset may hand out the same (stable) buffer every step
                    s_t.copy_(t, non_blocking=True)
            ent["graph"].replay()
            if ent.get("deferred"):
                opt._deferred = True          # the captured program stops short of the optimizer launch
                opt._deferred_lo = ent.get("deferred_lo")      # ... of all of it, or of the part it left to the merge
            self._graph_replays = getattr(self, "_graph_replays", 0) + 1
            ws["graph_launches"] = ws.get("graph_launches", 0) + ent["kernels"]
            return ent["out"]
        ent["calls"] += 1
        # two eager calls before the first capture of a job (allocations, lazy initialisation, library
        # heuristics); one is enough for a further signature of a job that already replays a graph - e.g. the short
        # last batch of an epoch, which comes round once per epoch and would otherwise run eagerly for two epochs
        warm = any(e.get("graph") is not None for k, e in cache.items() if k[0] == key[0])
        if ent["failed"] or ent["calls"] <= (1 if warm else 2):
            return self._eager_step(batch, idx, defer)
        try:
            # the captured program reads its inputs from fixed addresses: a private copy of each batch tensor,
            # or the tensor itself when the dataset marks it as a buffer it refills in place for every batch
            static = [t if getattr(t, "_kubeml_stable", False) else t.clone() for t in tensors]
            it = iter(static)
            sbatch = type(batch)(next(it) if isinstance(t, torch.Tensor) else t for t in batch) \
                if isinstance(batch, (tuple, list)) else static[0]
            torch.cuda.synchronize(self.device)
            k0 = _native_launches()
            g = torch.cuda.CUDAGraph()
            # thread_local: workers may be threads of one process, each capturing on its own GPU
            armed = bool(defer and opt.arm_deferral())
            try:
                with torch.cuda.graph(g, capture_error_mode="thread_local"):
                    out = ent["owner_train"](sbatch, idx) if "owner_train" in ent else self.train(sbatch, idx)
            finally:
                if armed:
                    opt.disarm()
            deferred = armed and bool(getattr(opt, "_deferred", False))
            ent.update(graph=g, static=static, out=out, owner=self, kernels=_native_launches() - k0,
                       deferred=deferred, deferred_lo=getattr(opt, "_deferred_lo", None) if deferred else None)
            g.replay()            # the capture itself did not execute: run this batch now
            ws["graph_launches"] = ws.get("graph_launches", 0) + ent["kernels"]
            return out
        except Exception as e:  # noqa: BLE001 - fall back to eager for this signature
            ent["failed"] = True
            self.logger.warning("CUDA graph capture of train() failed (%s); running eagerly", e)
            torch.cuda.synchronize(self.device)
            if getattr(opt, "_deferred", False):
                opt._deferred = False         # the aborted capture launched nothing
            return self._eager_step(batch, idx, False)

    # ---- task: val (network.py:312-360) ----------------------------------------------------------
    def _on_validation_start(self) -> None:
        self._set_device()
        self._network.eval()
        self._sync.attach(self)

    def __validate(self) -> Tuple[float, float, int]:
        self._on_validation_start()
        a = self.args
        assigned = split_minibatches(range(self._dataset.num_val_docs), a._N)[a._func_id]
        if len(assigned) == 0:
            return 0.0, 0.0, 0
        self._dataset._want_pinned = (self.device.type == "cuda"
                                      and self._dataset._has_batch_transform())
        self._dataset._load_validation_data(start=assigned.start, end=assigned.stop)
        batches = self._dataset._batches(self.batch_size, self.device)

        acc: Number = 0.0
        loss: Number = 0.0
        self._sync.load_reference(a._job_id, self._network)
        with torch.no_grad():
            for idx, batch in enumerate(batches):
                if not isinstance(batches, _sliced_type()):
                    batch = self._batch_to_device(batch)
                _acc, _loss = self.validate(batch, idx)
                acc = acc + _acc
                loss = loss + _loss
        n = len(batches)
        return float(acc) / n, float(loss) / n, len(self._dataset)

    # ---- task: infer (made coherent — load the reference model, call infer(data)) ----
    def __infer(self):
        data_json = self._inv.body
        if not data_json:
            self.logger.error("JSON not found in request")
            raise DataError()
        self._set_device()
        self._network.eval()
        try:
            self._sync.load_reference(self.args._job_id, self._network)
        except (KubeMLException, KeyError):
            self.logger.debug("no stored reference model for %s; using the function's weights",
                              self.args._job_id)
        with torch.no_grad():
            preds = self.infer(data_json)
        if isinstance(preds, torch.Tensor):
            return preds.cpu().numpy().tolist()
        if isinstance(preds, np.ndarray):
            return preds.tolist()
        if isinstance(preds, list):
            return preds
        raise InvalidFormatError()

    # ---- device (network.py:379-393) -----------------------------------------------------------
    def _set_device(self) -> None:
        if self.platform == "cpu" or not torch.cuda.is_available():
            self.device = torch.device("cpu")
            return
        gpu_id = get_gpu(self.args._func_id)
        self.device = torch.device(f"cuda:{gpu_id}")
        torch.cuda.set_device(self.device)
        self._network = self._network.to(self.device)
        self.logger.debug("Set device to %s", self.device)

    # ---- user overridables (network.py:463-476) ------------------------------------------------
    def configure_optimizers(self) -> torch.optim.Optimizer:
        pass

    def init(self):
        pass

    def train(self, batch, batch_index: int) -> Number:
        pass

    def validate(self, batch, batch_index: int) -> Tuple[Number, Number]:
        pass

    def infer(self, data: List[Any]) -> Union[torch.Tensor, np.ndarray, List[float]]:
        pass


class _LossStage:
    """Page-locked landing zone (a ring) for per-step scalars (loss): written asynchronously now, read after a
    sync.  The scalar travels by a one-thread kernel storing into the mapped host buffer (`kml_post_f32`), not by
    a copy-engine transfer, which would cost two engine hand-overs between consecutive steps
    (`KUBEML_LOSS_POST=copy` restores the memcpy)."""

    def __init__(self, device: torch.device, capacity: int = 4096):
        self.device = device
        self.buf = torch.zeros(capacity, dtype=torch.float32).pin_memory()
        self.cap = capacity
        self.head = 0           # first slot not read yet
        self.tail = 0           # next slot to write
        self.carry = 0.0
        self.bytes = 0
        self._kernel_store = os.environ.get("KUBEML_LOSS_POST", "kernel") != "copy"

    @property
    def n(self) -> int:
        return self.tail - self.head

    def push(self, t: torch.Tensor) -> None:
        if self.n == self.cap:
            self.carry += self.drain()
        slot = self.tail % self.cap
        dst = self.buf[slot:slot + 1]
        src = t.detach().reshape(1)
        if self._kernel_store and src.dtype == torch.float32 and src.is_contiguous():
            from .ops import nn as _K
            _K.post_f32(src, dst)
        else:
            dst.copy_(src, non_blocking=True)
        self.tail += 1
        self.bytes += 4

    def drain(self, count: Optional[int] = None, event=None) -> float:
        """Sum of the oldest `count` entries (all by default) once `event` (default: everything queued on the
        current stream) has happened."""
        n = self.n if count is None else min(int(count), self.n)
        v, self.carry = self.carry, 0.0
        if n == 0:
            return v
        if event is not None:
            event.synchronize()
        else:
            torch.cuda.current_stream(self.device).synchronize()
        lo, hi = self.head % self.cap, (self.head + n - 1) % self.cap + 1
        if lo < hi:
            v += float(self.buf[lo:hi].sum())
        else:
            v += float(self.buf[lo:].sum()) + float(self.buf[:hi].sum())
        self.head += n
        return v


def _fault_hook(spec):
    """`inject_fault` option: a callable (func_id, epoch, round, batch) -> bool, or the picklable text form
    "func@epoch[:round[:batch]][/stall=seconds]" (e.g. "1@2:1" = function 1 fails in epoch 2 at the start of
    round 1; with /stall=3 it sleeps 3 s there instead of failing) that
    process-per-GPU workers and the `--inject-fault` CLI flag use."""
    if spec is None or callable(spec):
        return spec
    try:
        text, _, stall = str(spec).partition("/stall=")
        who, _, when = text.partition("@")
        parts = [int(v) for v in when.split(":")] if when else []
        func, epoch = int(who), (parts[0] if parts else 1)
        rnd = parts[1] if len(parts) > 1 else 0
        bat = parts[2] if len(parts) > 2 else 0
        stall_s = float(stall) if stall else 0.0
    except ValueError:
        raise KubeMLException(f"bad inject_fault spec {spec!r} (want func@epoch[:round[:batch]][/stall=seconds])", 400)

    def hook(f, e, r, b):
        if not (f == func and int(e) == epoch and r == rnd and b == bat):
            return False
        if stall_s > 0:          # a straggler, not a failure: the worker goes quiet for a while
            import time
            time.sleep(stall_s)
            return False
        return True
    return hook


def _native_launches() -> int:
    try:
        from . import _native
        return _native.launch_count() if _native._lib is not None else 0
    except Exception:  # noqa: BLE001
        return 0


def _to(e, device):
    return e.to(device, non_blocking=True) if isinstance(e, torch.Tensor) else e


def _sliced_type():
    from .dataset import SlicedBatches
    return SlicedBatches
