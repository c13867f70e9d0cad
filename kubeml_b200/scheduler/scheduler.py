"""Scheduler (reference: ml/pkg/scheduler/scheduler.go:48-117, api.go:47-203).

`submit_train_task` mints the 8-hex job id and queues the task (api.go:78-116); a loop pops
tasks, asks the policy for a parallelism and tells the parameter server to start a new job or
update a running one (scheduler.go:48-89).  `update_job` is what a running train job calls
between epochs (POST /job, api.go:47-76).  `infer` runs the task's function with `task=infer`
against the job's reference model — the reference hard-codes the function name `network`
(api.go:140); here the request names the job, whose function is looked up in the PS.
"""
from __future__ import annotations

import logging
import threading
from typing import Any, Optional

from ..api.types import InferRequest, JobInfo, JobState, TrainRequest, TrainTask
from ..exceptions import KubeMLException
from ..httpd import Request, Router, Server
from ..util import create_job_id
from .policy import SchedulerPolicy, TaskOperation, ThroughputBasedPolicy
from .queue import SchedulerQueue


class Scheduler:
    def __init__(self, ps: Any = None, policy: Optional[SchedulerPolicy] = None,
                 max_parallelism: int = 8):
        self.ps = ps          # needs start_task(task), update_task(task), infer(job_id, data)
        self.policy = policy or ThroughputBasedPolicy(max_parallelism=max_parallelism)
        self.queue = SchedulerQueue()
        self.log = logging.getLogger("kubeml.scheduler")
        self._thread: Optional[threading.Thread] = None
        self._running = False

    # ---- API ---------------------------------------------------------------------------
    def submit_train_task(self, req: TrainRequest) -> str:
        job_id = create_job_id()
        task = TrainTask(request=req, job=JobInfo(id=job_id, state=JobState()))
        self.queue.push(task)
        return job_id

    def update_job(self, task: TrainTask) -> None:
        self.queue.push(task)

    def finish_job(self, task_id: str) -> None:
        """DELETE /finish/{taskId} (api.go:160-176): forget the policy's cache entry."""
        self.policy.task_finished(task_id)

    def infer(self, req: InferRequest) -> Any:
        if self.ps is None:
            raise KubeMLException("no parameter server attached", 503)
        return self.ps.infer(req.model_id, req.data)

    # ---- loop --------------------------------------------------------------------------
    def schedule_one(self, task: TrainTask) -> None:
        parallelism, op = self.policy.calculate_parallelism(task)
        task.job.state.parallelism = parallelism
        if self.ps is None:
            return
        try:
            if op is TaskOperation.CREATE:
                self.ps.start_task(task)
            else:
                self.ps.update_task(task)
        except Exception as e:  # noqa: BLE001 - scheduler.go:71-85 logs and moves on
            self.log.error("error scheduling task %s: %s", task.job.id, e)

    def _loop(self) -> None:
        while self._running:
            task = self.queue.pop(timeout=0.5)
            if task is not None:
                self.schedule_one(task)

    def start(self) -> "Scheduler":
        if self._thread is None:
            self._running = True
            self._thread = threading.Thread(target=self._loop, daemon=True, name="scheduler")
            self._thread.start()
        return self

    def stop(self) -> None:
        self._running = False
        self.queue.close()
        if self._thread is not None:
            self._thread.join(timeout=2)
            self._thread = None

    # ---- HTTP (scheduler/api.go:184-192) -------------------------------------------------------
    def router(self) -> Router:
        r = Router()
        r.add("GET", "/health", lambda q: (200, None))
        r.add("POST", "/train", lambda q: (200, self.submit_train_task(TrainRequest.from_dict(q.json())),
                                           "text/plain"))
        r.add("POST", "/job", lambda q: self.update_job(TrainTask.from_dict(q.json())))
        r.add("POST", "/infer", lambda q: self.infer(InferRequest.from_dict(q.json())))
        r.add("DELETE", "/finish/{taskId}", lambda q: self.finish_job(q.params["taskId"]))
        return r

    def serve(self, host: str, port: int) -> Server:
        return Server(self.router(), host, port).start()
