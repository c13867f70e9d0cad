"""A train job as a process of its own: `python -m kubeml_b200 --jobPort P --jobId ID`.

Reference: with STANDALONE_JOBS the parameter server creates a Pod + Service `job-<id>` running the same binary
with `--jobPort 9090 --jobId <id>` (ml/cmd/ml/main.go:60-156, ml/pkg/ps/job_pod.go:18-242); the pod serves
`/start`, `/update`, `/next/{funcId}`, `/stop`, `/health` (ml/pkg/train/api.go:30-149), is handed its task by the
parameter server (ml/pkg/ps/api.go:139-222), reports metrics and its end to the parameter server and asks the
scheduler for its next parallelism (ml/pkg/train/util.go, ml/pkg/scheduler/client/client.go) - all over HTTP.

Here the "pod" is a child process of the parameter server on the same host.  It shares `$KUBEML_HOME` (datasets,
functions, histories, checkpoints) with the control plane, owns its GPU workers (one thread per visible GPU, the
in-process `Cluster` machinery without the controller / scheduler loops) and talks back through the two small
HTTP proxies below.  What the isolation buys is what it buys the reference: a job that crashes the interpreter
or wedges a GPU context takes its own process with it, and the parameter server notices (`ps/parameter_server.py`).
"""
from __future__ import annotations

import logging
import os
import threading
from typing import Any, Optional

from ..api.types import JobState, MetricUpdate, TrainTask
from ..exceptions import KubeMLException
from ..httpd import Request, Router, Server, http_call


class PSProxy:
    """What a `TrainJob` calls on its parameter server, over HTTP (ml/pkg/ps/client/client.go:33-160)."""

    def __init__(self, url: str):
        self.url = url.rstrip("/")

    def update_metrics(self, job_id: str, m: MetricUpdate) -> None:
        http_call("POST", f"{self.url}/metrics/{job_id}", m.to_dict())

    def job_finished(self, job_id: str, err: Optional[BaseException] = None) -> None:
        body = None
        if err is not None:
            body = {"error": str(err), "code": int(getattr(err, "status_code", 500) or 500)}
        http_call("POST", f"{self.url}/finish/{job_id}", body or {})


class SchedulerProxy:
    """`UpdateJob` of ml/pkg/scheduler/client/client.go:36-122: the job asks for its next parallelism."""

    def __init__(self, url: str):
        self.url = url.rstrip("/")

    def update_job(self, task: TrainTask) -> None:
        http_call("POST", f"{self.url}/job", task.to_dict())


class JobServer:
    def __init__(self, job_id: str, home: Optional[str] = None, ps_url: Optional[str] = None,
                 scheduler_url: Optional[str] = None, options: Optional[dict] = None):
        from ..controller.cluster import Cluster
        self.job_id = job_id
        self.log = logging.getLogger(f"kubeml.job.{job_id}")
        # stores, function registry and the worker machinery of an in-process cluster; its own scheduler and
        # parameter server stay idle (start=False): this process answers to the ones that spawned it
        self.cluster = Cluster(home=home, options=options, start=False)
        self.ps = PSProxy(ps_url) if ps_url else None
        self.scheduler = SchedulerProxy(scheduler_url) if scheduler_url else None
        self.job = None
        self.done = threading.Event()
        self._mu = threading.Lock()
        self._server: Optional[Server] = None

    # ---- routes (train/api.go:141-149) -------------------------------------------------------------
    def router(self) -> Router:
        r = Router()
        r.add("GET", "/health", lambda q: (200, None))
        r.add("POST", "/start", lambda q: self.start_task(TrainTask.from_dict(q.json())))
        r.add("POST", "/update", lambda q: self.update_task(JobState.from_dict(q.json())))
        r.add("POST", "/next/{funcId}", self._http_next)
        r.add("DELETE", "/stop", lambda q: self.stop())
        return r

    def start_task(self, task: TrainTask) -> None:
        """POST /start (train/api.go:30-66); a second task for the same process is refused (the reference
        leaves that as a TODO)."""
        from .job import TrainJob
        if task.job.id != self.job_id:
            raise KubeMLException(f"this process serves job {self.job_id}, not {task.job.id}", 400)
        with self._mu:
            if self.job is not None:
                raise KubeMLException(f"job {self.job_id} already has its task", 400)
            c = self.cluster
            invoker = c.invoker_factory(c, task)
            sv = getattr(invoker, "services", None)
            device_mode = (getattr(sv, "device_sync", None) is not None
                           or "sync_factory" in (getattr(sv, "options", None) or {}))
            self.job = TrainJob(task, invoker, c.tensor_store, history_store=c.histories,
                                scheduler=self.scheduler, ps=self.ps, device_mode=device_mode,
                                checkpoint_dir=c.checkpoint_dir)
            self.job.trace_dir = os.path.join(c.home, "traces")
            c.jobs[task.job.id] = self.job          # the workers' merge barrier looks the job up here
        threading.Thread(target=self._run, daemon=True, name=f"job-{self.job_id}").start()

    def _run(self) -> None:
        try:
            self.job.train()            # reports to the parameter server itself (TrainJob._cleanup)
        finally:
            self.done.set()

    def update_task(self, state: JobState) -> None:
        if self.job is None:
            raise KubeMLException("no task yet", 404)
        self.job.update_task(state)

    def _http_next(self, q: Request) -> None:
        """POST /next/{funcId} (train/api.go:100-126): a function has published its replica and waits for the
        merge.  The workers of this process arrive by a direct call; the route serves functions that run
        elsewhere (`runtime/server.py`)."""
        if self.job is None:
            raise KubeMLException("no task yet", 404)
        self.job.next_iteration(self.job_id, int(q.params["funcId"]))

    def stop(self) -> None:
        if self.job is None:
            raise KubeMLException("no task yet", 404)
        self.job.stop()

    # ---- life cycle --------------------------------------------------------------------------------
    def serve(self, host: str, port: int) -> Server:
        self._server = Server(self.router(), host, port).start()
        return self._server

    def close(self) -> None:
        if self._server is not None:
            self._server.stop()
        self.cluster.shutdown()


class JobClient:
    """ml/pkg/train/client/client.go:31-107."""

    def __init__(self, url: str):
        self.url = url.rstrip("/")

    def health(self) -> bool:
        try:
            http_call("GET", self.url + "/health", timeout=2.0)
            return True
        except Exception:  # noqa: BLE001 - not up yet, or gone
            return False

    def start(self, task: TrainTask) -> None:
        http_call("POST", self.url + "/start", task.to_dict())

    def update(self, state: JobState) -> None:
        http_call("POST", self.url + "/update", state.to_dict())

    def stop(self) -> None:
        http_call("DELETE", self.url + "/stop")

    def next_iteration(self, func_id: int) -> None:
        http_call("POST", f"{self.url}/next/{int(func_id)}")


def main(job_id: str, port: int, host: str = "127.0.0.1") -> int:
    """Entry of the job process: serve until the task it was given has ended, then exit (the pod completes)."""
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(levelname)s %(name)s: %(message)s")
    import json
    srv = JobServer(job_id, ps_url=os.environ.get("KUBEML_PS_URL"), scheduler_url=os.environ.get("KUBEML_SCHEDULER_URL"),
                    options=json.loads(os.environ.get("KUBEML_JOB_OPTIONS") or "{}"))
    srv.serve(host, port)
    logging.getLogger("kubeml.job").info("job %s serving on %s:%d", job_id, host, port)
    idle = float(os.environ.get("KUBEML_JOB_IDLE_TIMEOUT_S", "120"))
    try:
        # a process that is never handed a task (its parameter server died) does not linger
        waited = 0.0
        while srv.job is None and waited < idle:
            if srv.done.wait(0.25):
                break
            waited += 0.25
        if srv.job is not None:
            srv.done.wait()
    except KeyboardInterrupt:
        if srv.job is not None:
            srv.job.stop()
            srv.done.wait(30)
    finally:
        srv.close()
    err = srv.job.exit_err if srv.job is not None else RuntimeError("no task received")
    return 0 if err is None else 1
