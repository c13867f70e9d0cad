"""`Cluster`: the whole control plane assembled in one process.

The reference deploys controller, scheduler, parameter server and storage as four Deployments
plus one pod per job (ml/charts/kubeml/templates/deployment.yaml:11-218) and the closest thing it
has to a single-process harness is ml/tests/integration.go:13-35, which starts three roles as
goroutines on debug ports.  On a single 8xB200 host that harness *is* the deployment: the roles
are plain objects wired by reference, optionally also served over HTTP (`serve()`), and jobs are
threads that drive warm workers.
"""
from __future__ import annotations

import logging
import os
import threading
from typing import Any, Callable, Dict, List, Optional

from ..api import const
from ..api.types import TrainTask
from ..exceptions import KubeMLException
from ..model.store import TensorStore
from ..ps.parameter_server import ParameterServer
from ..runtime.functions import FunctionRegistry
from ..runtime.services import Services
from ..scheduler.scheduler import Scheduler
from ..storage.service import StorageService
from ..storage.store import ShardStore
from ..train.history import HistoryStore
from ..train.invoker import Invoker, ThreadInvoker
from ..train.job import TrainJob
from .controller import Controller


class Cluster:
    def __init__(self, home: Optional[str] = None, max_parallelism: Optional[int] = None,
                 invoker_factory: Optional[Callable[["Cluster", TrainTask], Invoker]] = None,
                 options: Optional[dict] = None, start: bool = True, tensor_store: Optional[TensorStore] = None):
        self.home = home or const.home()
        os.makedirs(self.home, exist_ok=True)
        self.shard_store = ShardStore(os.path.join(self.home, "datasets"))
        # the store the workers publish into: with process-per-GPU workers it is the runtime's (runtime/spmd.py)
        self.tensor_store = tensor_store if tensor_store is not None else TensorStore()
        self.histories = HistoryStore(os.path.join(self.home, "history"))
        self.functions = FunctionRegistry(os.path.join(self.home, "functions"))
        self.checkpoint_dir = os.path.join(self.home, "checkpoints")
        self.options = dict(options or {})
        self.max_parallelism = max_parallelism or self._detect_max_parallelism()
        self.invoker_factory = invoker_factory or self._default_invoker

        self.scheduler = Scheduler(max_parallelism=self.max_parallelism)
        self.ps = ParameterServer(self._make_job, scheduler=self.scheduler)
        self.ps.job_options = self.options
        self.ps.finished_infer = self.infer_finished
        self._serving: Dict[str, Any] = {}          # task id -> (worker, request) of checkpoints being served
        self._serving_lock = threading.Lock()
        self.scheduler.ps = self.ps
        self.storage = StorageService(self.shard_store)
        self.controller = Controller(self.scheduler, self.ps, self.storage, self.histories,
                                     self.functions)
        self._servers: list = []
        self.jobs: Dict[str, TrainJob] = {}
        if start:
            self.scheduler.start()

    @staticmethod
    def _detect_max_parallelism() -> int:
        try:
            import torch
            if torch.cuda.is_available():
                return max(1, min(const.MAX_GPUS, torch.cuda.device_count()))
        except Exception:  # noqa: BLE001
            pass
        return const.MAX_GPUS

    # ---- job factory ---------------------------------------------------------------------
    def _default_invoker(self, cluster: "Cluster", task: TrainTask) -> Invoker:
        job_holder: Dict[str, Any] = {}
        services = Services(shard_store=self.shard_store, tensor_store=self.tensor_store,
                            barrier=lambda job_id, func_id: self.jobs[job_id].next_iteration(job_id, func_id),
                            options=self.options)
        device_sync = None
        try:
            import torch
            if torch.cuda.is_available() and not self.options.get("force_host_sync"):
                from ..parallel.device_sync import ThreadedDeviceSyncGroup
                device_sync = ThreadedDeviceSyncGroup.for_job(task.job.id, self.max_parallelism)
        except ImportError:
            device_sync = None
        if device_sync is not None:
            services.device_sync = device_sync
        gpu_ids = None
        try:
            import torch
            if torch.cuda.is_available():
                gpu_ids = list(range(torch.cuda.device_count()))
        except Exception:  # noqa: BLE001
            pass
        return ThreadInvoker(services, self.functions, max_parallelism=self.max_parallelism,
                             gpu_ids=gpu_ids)

    def _make_job(self, task: TrainTask, ps: ParameterServer) -> TrainJob:
        invoker = self.invoker_factory(self, task)
        sv = getattr(invoker, "services", None)
        # the workers merge among themselves (device kernel, or a collective on a GPU-less box): no host-side merge
        device_mode = (getattr(sv, "device_sync", None) is not None
                       or "sync_factory" in (getattr(sv, "options", None) or {}))
        job = TrainJob(task, invoker, self.tensor_store, history_store=self.histories,
                       scheduler=self.scheduler, ps=ps, device_mode=device_mode,
                       checkpoint_dir=self.checkpoint_dir)
        job.trace_dir = os.path.join(self.home, "traces")     # host timeline always; device spans with options["trace"]
        self.jobs[task.job.id] = job
        return job

    # ---- serving a finished task ---------------------------------------------------------------
    SERVING_SLOTS = 4

    def infer_finished(self, model_id: str, data: Any) -> Any:
        """`kubeml infer --id <task>` after the task has ended.  The reference can only ask a function named
        `network` about a model whose tensors its job deleted on exit (ml/pkg/scheduler/api.go:140,
        ml/pkg/train/util.go:208-244); here every finished task leaves a checkpoint, and inference
        runs the task's own function (`task=infer`) on it: the checkpoint is published once as the reference
        model `{task}:{layer}` of a private tensor store, a warm worker answers the requests."""
        from ..runtime.worker import Worker
        from ..utils.checkpoint import restore_into_store
        with self._serving_lock:
            ent = self._serving.get(model_id)
            if ent is None:
                try:
                    hist = self.histories.get(model_id)
                except Exception:  # noqa: BLE001 - unknown id
                    raise KubeMLException(f"model {model_id} not found", 404)
                if not os.path.isfile(os.path.join(self.checkpoint_dir, model_id, "model.pt")):
                    raise KubeMLException(f"task {model_id} left no checkpoint (it failed or was stopped)", 404)
                store = TensorStore()
                restore_into_store(self.checkpoint_dir, model_id, store)
                services = Services(shard_store=self.shard_store, tensor_store=store, barrier=None, options=self.options)
                worker = Worker(services, self.functions, gpu_id=0 if self._cuda() else None, name=f"serve-{model_id}")
                while len(self._serving) >= self.SERVING_SLOTS:            # oldest model makes room
                    self._serving.pop(next(iter(self._serving)))
                ent = self._serving[model_id] = (worker, hist.task, threading.Lock())
        worker, req, lock = ent
        q = {"task": "infer", "jobId": model_id, "N": 1, "K": -1, "funcId": 0, "batchSize": req.batch_size,
             "lr": req.lr, "epoch": 1}
        with lock:                                   # one request at a time per model: the worker owns one stream
            payload, status = worker.invoke(req.function_name, q, body=data)
        if status != 200:
            raise KubeMLException(str(payload.get("error", payload)), status)
        return payload.get("predictions")

    @staticmethod
    def _cuda() -> bool:
        try:
            import torch
            return torch.cuda.is_available()
        except Exception:  # noqa: BLE001
            return False

    # ---- convenience ---------------------------------------------------------------------
    def wait(self, job_id: str, timeout: Optional[float] = None) -> Optional[str]:
        """Block until the job has ended; its error text or None.  (The parameter server waits for the scheduler
        thread to have created the job first - as a thread of this process or, with STANDALONE_JOBS, as a
        process of its own, which never appears in `self.jobs`.)"""
        return self.ps.wait(job_id, timeout)

    def serve(self, host: str = const.HOST, controller_port: int = const.CONTROLLER_PORT,
              scheduler_port: Optional[int] = None, ps_port: Optional[int] = None,
              storage_port: Optional[int] = None) -> "Cluster":
        self._servers.append(self.controller.serve(host, controller_port))
        if scheduler_port:
            self._servers.append(self.scheduler.serve(host, scheduler_port))
            self.ps.scheduler_url = f"http://{'127.0.0.1' if host in ('0.0.0.0', '') else host}:{self._servers[-1].port}"
        if ps_port:
            self._servers.append(self.ps.serve(host, ps_port))
        if storage_port:
            self._servers.append(self.storage.serve(host, storage_port))
        return self

    @property
    def controller_port(self) -> int:
        return self._servers[0].port if self._servers else 0

    def shutdown(self) -> None:
        self.scheduler.stop()
        self.ps.close()
        for s in self._servers:
            s.stop()
        self._servers.clear()
