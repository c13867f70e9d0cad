"""Train job: owns one training task from init to history (reference: ml/pkg/train/job.go,
function.go, util.go, api.go).

Lifecycle (job.go:156-265): init -> for each epoch: fan the function out to N workers and merge
their replicas every K local steps -> (dynamic parallelism) ask the scheduler for next epoch's N
-> validate on the configured cadence -> stop on request or on reaching the goal accuracy -> final
validation -> persist the History -> report to the parameter server.

Design differences from the reference:
* functions run on warm workers through an `Invoker`, not through HTTP;
* on the host path replicas are merged by `MergeBarrier`; on the device path (`services.device_sync`)
  the merge is a kernel launched by the workers themselves and the job only keeps books;
* the final reference model is kept (and can be checkpointed) instead of being deleted with the
  replicas (train/util.go:218 deletes both);
* parallelism is clamped to [1, invoker.max_parallelism].
"""
from __future__ import annotations

import logging
import threading
import time
from typing import Any, Callable, Dict, List, Optional

from ..api.types import (History, JobHistory, JobState, MetricUpdate, TrainTask, latest_metrics)
from ..exceptions import KubeMLException, MergeError
from ..model.model import Model, ParallelSGD
from ..model.store import TensorStore
from ..util import is_debug_env, limit_parallelism
from .barrier import MergeBarrier
from .history import HistoryStore
from .invoker import Invoker, Result


# Fan-outs of a job id, counted for the life of the control-plane process: the device merge derives its
# barrier generations from it (kubeml_b200.parallel.device_sync.RankSync._gen), and those must never
# repeat while the workers keep the job's arena - also when a new TrainJob object reuses a job id.
_SYNC_SEQ: Dict[str, int] = {}
_SYNC_SEQ_LOCK = threading.Lock()


def _next_sync_seq(job_id: str) -> int:
    with _SYNC_SEQ_LOCK:
        _SYNC_SEQ[job_id] = _SYNC_SEQ.get(job_id, 0) + 1
        return _SYNC_SEQ[job_id]


class TrainJob:
    def __init__(self, task: TrainTask, invoker: Invoker, tensor_store: TensorStore,
                 history_store: Optional[HistoryStore] = None,
                 scheduler: Any = None, ps: Any = None,
                 logger: Optional[logging.Logger] = None,
                 device_mode: bool = False,
                 checkpoint_dir: Optional[str] = None,
                 skip_final_validation: bool = False):
        self.task = task
        self.job_id = task.job.id
        self.invoker = invoker
        self.tensor_store = tensor_store
        self.history_store = history_store
        self.scheduler = scheduler      # needs .update_job(task)   (scheduler/client UpdateJob)
        self.ps = ps                    # needs .update_metrics(id, MetricUpdate), .job_finished(id, err)
        self.log = logger or logging.getLogger(f"kubeml.job.{self.job_id}")
        self.device_mode = device_mode
        self.checkpoint_dir = checkpoint_dir
        self.skip_final_validation = skip_final_validation
        self.release_on_finish = True

        opts = task.request.options
        self.parallelism = self._clamp(task.job.state.parallelism or opts.default_parallelism or 1)
        self.static = bool(opts.static_parallelism)
        self.validate_every = int(opts.validate_every)
        self.K = int(opts.k)
        self.goal_accuracy = float(opts.goal_accuracy)
        self.epoch = 0
        self.history = JobHistory()
        self.model: Optional[Model] = None
        self.optimizer = ParallelSGD()
        self.exit_err: Optional[BaseException] = None
        self.accuracy_reached = False
        self.start_time = 0.0
        self.finished = threading.Event()

        self._stop = threading.Event()
        self._sched_update: "list[JobState]" = []
        self._sched_cv = threading.Condition()
        self._barrier: Optional[MergeBarrier] = None
        self._barrier_lock = threading.Lock()
        # Timeline of the job: host spans of every phase, and - when the workers were asked to trace
        # (`options["trace"]` / KUBEML_TRACE=1) - their device spans (steps, H2D copies, merges) on one row per
        # worker and stream; written to <trace_dir>/<job>.trace.json for chrome://tracing / Perfetto.
        from ..utils.tracing import ChromeTrace
        self.trace = ChromeTrace()
        self.worker_traces: dict = {}
        self.trace_dir: Optional[str] = None
        self.merge_log: List[Dict[str, Any]] = []   # (epoch, participants, seconds) per merge
        self.epoch_stats: List[Dict[str, Any]] = []  # per epoch: device-timed seconds, steps, bytes

    # ------------------------------------------------------------------ external API
    def next_iteration(self, job_id: str, func_id: int) -> None:
        """Barrier arrival of a worker (POST /next/{funcId}, train/api.go:100-126)."""
        with self._barrier_lock:
            b = self._barrier
        if b is None:
            raise MergeError(RuntimeError("no epoch in progress"))
        b.arrive(func_id)

    def update_task(self, state: JobState) -> None:
        """New parallelism from the scheduler (POST /update, train/api.go:69-96)."""
        with self._sched_cv:
            self._sched_update.append(state)
            self._sched_cv.notify_all()

    def stop(self) -> None:
        """DELETE /stop (train/api.go:129-134); honoured once per epoch (job.go:235-239)."""
        self._stop.set()

    # ------------------------------------------------------------------ main loop
    def train(self) -> JobHistory:
        self.log.info("Starting train job %s (function=%s dataset=%s N=%d K=%d)", self.job_id,
                      self.task.request.function_name, self.task.request.dataset,
                      self.parallelism, self.K)
        try:
            self._run()
        except BaseException as e:  # noqa: BLE001
            self.log.error("job failed: %s", e)
            self.exit_err = e
        finally:
            self._cleanup()
        return self.history

    def _run(self) -> None:
        with self.trace.span("init", job=self.job_id):
            self._init()
        self.start_time = time.time()
        req = self.task.request
        self.epoch = 1
        while self.epoch <= req.epochs:
            with self.trace.span(f"epoch {self.epoch}", parallelism=self.parallelism):
                self._train_epoch()

            if not self.static and self.epoch < req.epochs and self.scheduler is not None:
                self._ask_scheduler()

            if (self.validate_every != 0 and self.epoch % self.validate_every == 0
                    and self.epoch != req.epochs):
                try:
                    with self.trace.span(f"validate {self.epoch}"):
                        self._validate()
                except KubeMLException as e:
                    self.log.error("error performing validation: %s", e)

            if self._stop.is_set():
                self.accuracy_reached = True
                self.exit_err = KubeMLException("job was force stopped", 499)
                break
            if self.accuracy_reached:
                self.log.info("goal accuracy reached, exiting")
                break
            self.epoch += 1

        if not self.accuracy_reached and not self.skip_final_validation:
            try:
                self._validate()
            except KubeMLException as e:
                self.log.error("error performing validation: %s", e)
        self._save_history()
        self.log.info("Training finished after %d epochs", min(self.epoch, req.epochs))
        if self.trace_dir:
            try:
                import os
                os.makedirs(self.trace_dir, exist_ok=True)
                self.trace.save(os.path.join(self.trace_dir, f"{self.job_id}.trace.json"))
            except OSError as e:
                self.log.error("could not write the trace: %s", e)

    # ------------------------------------------------------------------ phases
    def _query(self, task: str, func_id: int, n: int) -> Dict[str, Any]:
        """Same parameter names as buildFunctionURL (function.go:44-68)."""
        r = self.task.request
        return {"task": task, "jobId": self.job_id, "N": n, "K": self.K, "funcId": func_id,
                "batchSize": r.batch_size, "lr": r.lr, "epoch": self.epoch}

    def _init(self) -> None:
        q = self._query("init", 0, 1)
        resume = getattr(self.task.request.options, "resume_from", "")
        if resume:
            if not self.checkpoint_dir:
                raise KubeMLException("resume_from needs a checkpoint directory", 400)
            import os
            if not os.path.isfile(os.path.join(self.checkpoint_dir, resume, "model.pt")):
                raise KubeMLException(f"no checkpoint of task {resume}", 404)
            q.update(resumeFrom=resume, ckptDir=self.checkpoint_dir)
            self.log.info("resuming from the checkpoint of task %s", resume)
        payload, status = self.invoker.invoke_one(0, self.task.request.function_name, q)
        if status != 200:
            raise KubeMLException(f"error invoking init function: {payload}", status)
        layers = list(payload or [])
        if not layers:
            raise KubeMLException("length of the layers is zero", 500)
        self.layer_names = layers
        if not self.device_mode:
            self.model = Model(self.job_id, layers, self.tensor_store,
                               name=self.task.request.model_type, logger=self.log)
            self.model.build()
            self.model.summary()

    def _train_epoch(self) -> None:
        n = self.parallelism
        self.log.info("Started new epoch %d with parallelism %d", self.epoch, n)
        barrier = None
        if not self.device_mode:
            barrier = MergeBarrier(self.model, self.optimizer, n, on_merge=self._merged)
            with self._barrier_lock:
                self._barrier = barrier

        start = time.time()
        seq = _next_sync_seq(self.job_id)
        queries = [dict(self._query("train", i, n), syncSeq=seq) for i in range(n)]
        on_done = (lambda fid, ok: barrier.finished(fid, ok)) if barrier is not None else None
        results = self.invoker.invoke_many(self.task.request.function_name, queries, on_done)
        ok = self._check_function_errors(results)
        if barrier is not None:
            if barrier.failed is not None:
                raise MergeError(barrier.failed)
            barrier.done.wait()
            with self._barrier_lock:
                self._barrier = None
        elapsed = time.time() - start

        # unweighted mean of the per-function mean losses (train/util.go:82-95)
        loss = sum(float(p["loss"]) for _, p, _ in ok) / len(ok)
        merge_s = (barrier.merge_seconds if barrier is not None
                   else max(float(p.get("merge_seconds", 0.0)) for _, p, _ in ok))
        self.task.job.state.elapsed_time = elapsed
        self.task.job.state.parallelism = n
        self.history.parallelism.append(float(n))
        self.history.epoch_duration.append(time.time() - self.start_time)   # cumulative (job.go:327)
        self.history.train_loss.append(loss)
        self.history.merge_seconds.append(merge_s)
        self.epoch_stats.append({
            "epoch": self.epoch, "parallelism": n, "wall_seconds": elapsed,
            "train_seconds": max(float(p.get("train_seconds", 0.0)) for _, p, _ in ok),
            "merge_seconds": merge_s,
            "steps": max(int(p.get("steps", 0)) for _, p, _ in ok),
            "h2d_bytes": sum(int(p.get("h2d_bytes", 0)) for _, p, _ in ok),
            "d2h_bytes": sum(int(p.get("d2h_bytes", 0)) for _, p, _ in ok),
            "gpu_launches": sum(int(p.get("gpu_launches", 0)) for _, p, _ in ok),
            "merges": max(int(p.get("merges", 0)) for _, p, _ in ok),
            "merge_wait_seconds": max(float(p.get("merge_wait_seconds", 0.0)) for _, p, _ in ok),
            "fused_last_steps": max(int(p.get("fused_last_steps", 0)) for _, p, _ in ok)})
        epoch_t0_us = (start - (time.time() - time.perf_counter()) - self.trace._t0) * 1e6
        for fid, p, _ in ok:
            tr = p.get("trace") if isinstance(p, dict) else None
            if not tr:
                continue
            self.worker_traces[fid] = tr        # raw device spans of the last traced epoch (bench.py's gap table)
            for lane, key in enumerate(("step", "h2d", "merge")):
                for r, b, t_ms, d_ms in tr.get(key, []):
                    self.trace.device_span(f"{key} r{r} b{b}" if key != "merge" else f"merge {r}",
                                           epoch_t0_us + 1e3 * t_ms, 1e3 * d_ms, tid=fid * 4 + lane,
                                           epoch=self.epoch, worker=fid, stream=key)
        samples = self._epoch_samples()
        self.history.images_per_second.append(samples / elapsed if elapsed > 0 and samples else 0.0)
        self._push_metrics()

    def _epoch_samples(self) -> int:
        try:
            from ..storage import default_store
            store = getattr(getattr(self.invoker, "services", None), "shard_store", None) or default_store()
            return int(store.summary(self.task.request.dataset).train_set_size)
        except Exception:  # noqa: BLE001 - throughput is an optional extension
            return 0

    def _merged(self, participants: List[int], seconds: float) -> None:
        self.merge_log.append({"epoch": self.epoch, "participants": participants,
                               "seconds": seconds})

    def _validate(self) -> None:
        n = self.parallelism
        queries = [self._query("val", i, n) for i in range(n)]
        results = self.invoker.invoke_many(self.task.request.function_name, queries, None)
        ok = self._check_function_errors(results)
        # length-weighted means (train/util.go:100-122)
        total = sum(float(p["length"]) for _, p, _ in ok)
        if total <= 0:
            raise KubeMLException("validation returned no datapoints", 500)
        acc = sum(float(p["accuracy"]) * float(p["length"]) for _, p, _ in ok) / total
        loss = sum(float(p["loss"]) * float(p["length"]) for _, p, _ in ok) / total
        self.history.validation_loss.append(loss)
        self.history.accuracy.append(acc)
        self._push_metrics()
        if acc >= self.goal_accuracy:
            self.accuracy_reached = True

    def _ask_scheduler(self) -> None:
        """job.go:195-215: send the task (with last epoch's elapsed time), block for the answer."""
        with self._sched_cv:
            self._sched_update.clear()
        try:
            self.scheduler.update_job(self.task)
        except Exception as e:  # noqa: BLE001 - the reference logs and carries on (job.go:197-202)
            self.log.error("Error updating parallelism: %s", e)
            return
        with self._sched_cv:
            if not self._sched_cv.wait_for(lambda: bool(self._sched_update), timeout=60.0):
                self.log.error("scheduler did not answer; keeping parallelism %d", self.parallelism)
                return
            update = self._sched_update.pop()
        self.task.job.state = JobState(parallelism=update.parallelism,
                                       elapsed_time=self.task.job.state.elapsed_time)
        if not is_debug_env() and not limit_parallelism():
            self.parallelism = self._clamp(update.parallelism)
            self.log.info("Received next config from the scheduler: parallelism %d", self.parallelism)

    # ------------------------------------------------------------------ helpers
    def _clamp(self, n: int) -> int:
        return max(1, min(int(n), int(getattr(self.invoker, "max_parallelism", n) or n)))

    def _check_function_errors(self, results: List[Result]) -> List[Result]:
        """An epoch stands if at least one function answered (train/util.go:144-166)."""
        ok = [r for r in results if r[2] is None and r[1] is not None]
        for fid, _, err in results:
            if err is not None:
                self.log.error("function %d failed: %s", fid, err)
        if not ok:
            first = next((e for _, _, e in results if e is not None), None)
            raise KubeMLException(f"error invoking functions: {first}", getattr(first, "status_code", 500))
        return ok

    def _push_metrics(self) -> None:
        if self.ps is None:
            return
        try:
            self.ps.update_metrics(self.job_id, latest_metrics(self.history))
        except Exception as e:  # noqa: BLE001
            self.log.error("error updating metrics: %s", e)

    def _save_history(self) -> None:
        if self.history_store is not None:
            self.history_store.save(History(id=self.job_id, task=self.task.request, data=self.history))

    def _cleanup(self) -> None:
        """Drop worker replicas, keep the reference model (fix of train/util.go:208-244), tell
        the parameter server (job.go:161-170)."""
        try:
            self.tensor_store.delete(self.tensor_store.keys(f"{self.job_id}:*/*"))
        except Exception:  # noqa: BLE001
            pass
        if self.checkpoint_dir and self.exit_err is None:
            try:
                if self.device_mode:       # rank 0 writes its arena; history goes next to it
                    q = dict(self._query("checkpoint", 0, 1), ckptDir=self.checkpoint_dir)
                    payload, status = self.invoker.invoke_one(0, self.task.request.function_name, q)
                    if status != 200:
                        raise KubeMLException(str(payload), status)
                    import json
                    import os
                    with open(os.path.join(self.checkpoint_dir, self.job_id, "history.json"), "w") as f:
                        json.dump(History(id=self.job_id, task=self.task.request, data=self.history).to_dict(), f)
                else:
                    from ..utils.checkpoint import save_checkpoint
                    save_checkpoint(self.checkpoint_dir, self.job_id, self.tensor_store,
                                    getattr(self, "layer_names", []), self.task.request, self.history)
            except Exception as e:  # noqa: BLE001
                self.log.error("checkpoint failed: %s", e)
        release = getattr(self.invoker, "release_job", None)
        if callable(release) and self.release_on_finish:
            try:
                release(self.job_id)
            except Exception as e:  # noqa: BLE001
                self.log.error("could not release worker state: %s", e)
        if self.ps is not None:
            try:
                self.ps.job_finished(self.job_id, self.exit_err)
            except Exception as e:  # noqa: BLE001
                self.log.error("could not notify the parameter server: %s", e)
        self.finished.set()
