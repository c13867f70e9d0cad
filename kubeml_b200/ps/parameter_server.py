"""Parameter server (reference: ml/pkg/ps/api.go:16-360, parameter_server.go:73-105,
job_pod.go:18-242).

The reference keeps `jobIndex[jobId] -> TrainTask`, and for every new task either creates a
Kubernetes Pod+Service `job-<id>` running the job binary or (STANDALONE_JOBS=false) runs the job
as a goroutine.  On one host the second form is the only one that makes sense: a job is a thread
of the control plane driving warm GPU workers through an invoker chosen by `job_factory`.
The retry-on-start loop (api.go:192-207) disappears with the pod.
"""
from __future__ import annotations

import logging
import threading
from typing import Any, Callable, Dict, List, Optional

from ..api.types import JobState, MetricUpdate, TrainTask
from ..exceptions import KubeMLException
from ..httpd import Request, Router, Server
from ..train.job import TrainJob
from .metrics import MetricsRegistry

JobFactory = Callable[[TrainTask, "ParameterServer"], TrainJob]


class _JobProcess:
    """Handle of a job that runs as a child process (`task.job.handle` in stand-alone mode): what the reference
    keeps as `JobInfo.Pod` / `JobInfo.Svc` (ml/pkg/api/types.go:59-65) plus the job client that talks to it."""

    def __init__(self, proc, client, url: str):
        self.proc, self.client, self.url = proc, client, url

    def update_task(self, state: JobState) -> None:
        self.client.update(state)

    def stop(self) -> None:
        self.client.stop()


class ParameterServer:
    def __init__(self, job_factory: JobFactory, scheduler: Any = None, standalone_jobs: Optional[bool] = None):
        self.job_factory = job_factory
        self.scheduler = scheduler
        # ml/cmd/ml/main.go:119-129: STANDALONE_JOBS picks pod-per-job over goroutine-per-job.  Here: a child
        # process per job (`train/server.py`) over a thread per job; the thread is the default on one host.
        if standalone_jobs is None:
            import os
            standalone_jobs = os.environ.get("STANDALONE_JOBS", "").lower() in ("1", "true", "yes")
        self.standalone_jobs = bool(standalone_jobs)
        self.url: Optional[str] = None              # where job processes reach this server (set by serve())
        self.scheduler_url: Optional[str] = None    # ... and the scheduler
        self._own_servers: list = []
        self.job_options: dict = {}                 # JSON-able worker options handed to job processes
        self.finished_infer: Optional[Callable[[str, Any], Any]] = None   # inference on a finished task's checkpoint
        self.metrics = MetricsRegistry()
        self.log = logging.getLogger("kubeml.ps")
        self._index: Dict[str, TrainTask] = {}
        self._threads: Dict[str, threading.Thread] = {}
        self._finished: Dict[str, Optional[str]] = {}    # job id -> error text (None = ok)
        self._mu = threading.Lock()
        self._log_handlers: Dict[str, logging.Handler] = {}

    # ---- scheduler-facing ----------------------------------------------------------------
    def start_task(self, task: TrainTask) -> None:
        """POST /start (api.go:139-222)."""
        with self._mu:
            if task.job.id in self._index:
                raise KubeMLException(f"task {task.job.id} already exists", 400)
            self._index[task.job.id] = task
        if self.standalone_jobs:
            try:
                self._start_job_process(task)
            except BaseException:
                with self._mu:
                    self._index.pop(task.job.id, None)      # api.go:170-176: no resources, no entry
                raise
            self.metrics.task_started("train")
            return
        self._open_job_log(task.job.id)
        job = self.job_factory(task, self)
        task.job.handle = job
        self.metrics.task_started("train")
        t = threading.Thread(target=job.train, daemon=True, name=f"job-{task.job.id}")
        with self._mu:
            self._threads[task.job.id] = t
            t.start()

    # ---- stand-alone jobs: one process per job (ps/job_pod.go:18-242, ps/api.go:162-207) -------------------
    START_RETRIES = 10          # api.go:192
    READY_TIMEOUT_S = 20.0      # job_pod.go:172-186 waits that long for the pod to be Running

    def _start_job_process(self, task: TrainTask) -> None:
        import os
        import subprocess
        import sys
        import time
        from ..api import const
        from ..train.server import JobClient
        from ..util import find_free_port
        job_id = task.job.id
        self._ensure_reachable()
        port = find_free_port()
        env = dict(os.environ)
        env["KUBEML_PS_URL"] = self.url
        if self.scheduler_url:
            env["KUBEML_SCHEDULER_URL"] = self.scheduler_url
        env["KUBEML_HOME"] = const.home()
        if self.job_options:
            import json
            env["KUBEML_JOB_OPTIONS"] = json.dumps({k: v for k, v in self.job_options.items()
                                                    if isinstance(v, (bool, int, float, str, type(None)))})
        env.pop("STANDALONE_JOBS", None)
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        env["PYTHONPATH"] = root + (os.pathsep + env["PYTHONPATH"] if env.get("PYTHONPATH") else "")
        log_dir = os.path.join(const.home(), "logs")
        os.makedirs(log_dir, exist_ok=True)
        # the child's output *is* the job log (`kubeml logs --id <job>` = `kubectl logs job-<id>`, cmd/log.go:28-65)
        logf = open(os.path.join(log_dir, f"job-{job_id}.log"), "ab", buffering=0)
        proc = subprocess.Popen([sys.executable, "-m", "kubeml_b200", "--jobPort", str(port), "--jobId", job_id],
                                env=env, stdout=logf, stderr=subprocess.STDOUT, stdin=subprocess.DEVNULL)
        logf.close()
        url = f"http://127.0.0.1:{port}"
        client = JobClient(url)
        handle = _JobProcess(proc, client, url)
        deadline = time.monotonic() + float(os.environ.get("KUBEML_JOB_READY_TIMEOUT_S", self.READY_TIMEOUT_S))
        while not client.health():                      # the readiness probe of job_pod.go:120-131
            if proc.poll() is not None or time.monotonic() > deadline:
                self._kill(proc)
                raise KubeMLException(f"job process of {job_id} did not become ready "
                                      f"(exit code {proc.poll()}, see logs/job-{job_id}.log)", 500)
            time.sleep(0.05)
        last = None
        for i in range(self.START_RETRIES):             # api.go:192-207: 10 tries, 200 ms x 2i apart
            try:
                client.start(task)
                last = None
                break
            except KubeMLException as e:
                last = e
                if e.status_code < 500 or proc.poll() is not None:
                    break
                self.log.debug("error sending the task to job %s, retrying: %s", job_id, e)
                time.sleep(0.2 * 2 * i)
        if last is not None:
            self._kill(proc)
            raise KubeMLException(f"unable to send task for job {job_id}: {last}", 500)
        task.job.handle = handle
        t = threading.Thread(target=self._watch_job_process, args=(job_id, proc), daemon=True, name=f"job-{job_id}")
        with self._mu:
            self._threads[job_id] = t
            t.start()

    def _watch_job_process(self, job_id: str, proc) -> None:
        """The job's liveness: a process that ends without having reported `/finish` (crash, kill, wedged GPU
        context torn down by the driver) is a failed job, not a job that runs for ever in `task list`."""
        rc = proc.wait()
        with self._mu:
            reported = job_id in self._finished
        if not reported:
            self.job_finished(job_id, KubeMLException(f"job process exited with code {rc} without reporting", 500))

    @staticmethod
    def _kill(proc) -> None:
        if proc.poll() is None:
            proc.terminate()
            try:
                proc.wait(timeout=5)
            except Exception:  # noqa: BLE001
                proc.kill()

    def _ensure_reachable(self) -> None:
        """Job processes call back over HTTP: make sure this server and the scheduler listen somewhere."""
        from ..util import find_free_port
        if self.url is None:
            srv = self.serve("127.0.0.1", find_free_port())
            self._own_servers.append(srv)
        if self.scheduler_url is None and self.scheduler is not None and hasattr(self.scheduler, "serve"):
            srv = self.scheduler.serve("127.0.0.1", find_free_port())
            self._own_servers.append(srv)
            self.scheduler_url = f"http://127.0.0.1:{srv.port}"

    def close(self) -> None:
        """Stop what stand-alone mode started: listeners and job processes still running."""
        with self._mu:
            handles = [t.job.handle for t in self._index.values() if isinstance(t.job.handle, _JobProcess)]
        for h in handles:
            self._kill(h.proc)
        for srv in self._own_servers:
            srv.stop()
        self._own_servers.clear()

    # ---- per-job log files: what `kubeml logs --id <job> [-f]` reads -------------------------------
    # The reference's `logs` verb is `kubectl logs job-<id>` (ml/pkg/kubeml-cli/cmd/log.go:28-65): the
    # job pod's stdout.  A job here is a thread, so its logger `kubeml.job.<id>` (the train job, and the
    # workers' `kubeml.job.<id>.worker*` children) gets a file handler for the lifetime of the job.
    def _open_job_log(self, job_id: str) -> None:
        import os
        from ..api import const
        log_dir = os.path.join(const.home(), "logs")
        try:
            os.makedirs(log_dir, exist_ok=True)
            h = logging.FileHandler(os.path.join(log_dir, f"job-{job_id}.log"), delay=False)
        except OSError as e:
            self.log.warning("cannot open the log file of job %s: %s", job_id, e)
            return
        h.setFormatter(logging.Formatter("%(asctime)s %(levelname)s %(name)s: %(message)s"))
        lg = logging.getLogger(f"kubeml.job.{job_id}")
        if lg.level == logging.NOTSET or lg.level > logging.INFO:
            lg.setLevel(logging.INFO)
        lg.addHandler(h)
        with self._mu:
            self._log_handlers[job_id] = h

    def _close_job_log(self, job_id: str) -> None:
        with self._mu:
            h = self._log_handlers.pop(job_id, None)
        if h is not None:
            logging.getLogger(f"kubeml.job.{job_id}").removeHandler(h)
            h.close()

    def update_task(self, task: TrainTask) -> None:
        """POST /update/{jobId} (api.go:72-119): relay the new parallelism to the job."""
        with self._mu:
            cur = self._index.get(task.job.id)
        if cur is None or cur.job.handle is None:
            raise KubeMLException(f"task {task.job.id} does not exist", 404)
        cur.job.handle.update_task(JobState(parallelism=task.job.state.parallelism,
                                            elapsed_time=task.job.state.elapsed_time))

    # ---- job-facing ----------------------------------------------------------------------
    def update_metrics(self, job_id: str, m: MetricUpdate) -> None:
        """POST /metrics/{jobId} (api.go:226-257)."""
        self.metrics.update(job_id, m)

    def job_finished(self, job_id: str, err: Optional[BaseException] = None) -> None:
        """POST /finish/{jobId} (api.go:261-333): clear gauges, tell the scheduler, drop the entry."""
        self.metrics.clear(job_id)
        self.metrics.task_finished("train")
        if self.scheduler is not None:
            try:
                self.scheduler.finish_job(job_id)
            except Exception as e:  # noqa: BLE001
                self.log.error("error notifying scheduler: %s", e)
        with self._mu:
            self._index.pop(job_id, None)
            self._finished[job_id] = None if err is None else str(err)
        if err is not None:
            self.log.error("job %s exited with error: %s", job_id, err)
            logging.getLogger(f"kubeml.job.{job_id}").error("job exited with error: %s", err)
        self._close_job_log(job_id)

    # ---- controller-facing ---------------------------------------------------------------
    def list_tasks(self) -> List[TrainTask]:
        with self._mu:
            return list(self._index.values())

    def stop_task(self, job_id: str) -> None:
        """DELETE /stop/{jobId} (api.go:16-66)."""
        with self._mu:
            task = self._index.get(job_id)
        if task is None or task.job.handle is None:
            raise KubeMLException(f"task {job_id} not found", 404)
        task.job.handle.stop()

    def wait(self, job_id: str, timeout: Optional[float] = None) -> Optional[str]:
        """Block until a job's thread ended; returns its error text or None (no reference
        counterpart; the CLI there polls `task list`)."""
        import time
        deadline = None if timeout is None else time.monotonic() + timeout
        while True:
            with self._mu:
                t = self._threads.get(job_id)
                known = t is not None or job_id in self._finished
            if known:
                break
            if deadline is not None and time.monotonic() > deadline:
                raise TimeoutError(job_id)
            time.sleep(0.005)
        if t is not None:
            t.join(None if deadline is None else max(0.0, deadline - time.monotonic()))
            if t.is_alive():
                raise TimeoutError(job_id)
        with self._mu:
            return self._finished.get(job_id)

    def infer(self, model_id: str, data: Any) -> Any:
        """Run `task=infer` of the job's function on its reference model."""
        with self._mu:
            task = self._index.get(model_id)
        if (task is None or task.job.handle is None) and self.finished_infer is not None:
            # the task has ended: serve its checkpoint (controller/cluster.py `infer_finished`)
            self.metrics.task_started("inference")
            try:
                return self.finished_infer(model_id, data)
            finally:
                self.metrics.task_finished("inference")
        if task is None or task.job.handle is None:
            raise KubeMLException(f"model {model_id} not found (is the job still running?)", 404)
        if isinstance(task.job.handle, _JobProcess):
            raise KubeMLException("the model lives in the job's own process (STANDALONE_JOBS); "
                                  "inference runs against jobs of this process", 501)
        job: TrainJob = task.job.handle
        self.metrics.task_started("inference")
        try:
            q = job._query("infer", 0, 1)
            payload, status = job.invoker.invoke_one(0, task.request.function_name, q, body=data)
            if status != 200:
                raise KubeMLException(str(payload.get("error", payload)), status)
            return payload.get("predictions")
        finally:
            self.metrics.task_finished("inference")

    # ---- HTTP (ps/api.go:335-345 + :8080/metrics) ------------------------------------------------
    def router(self) -> Router:
        r = Router()
        r.add("GET", "/health", lambda q: (200, None))
        r.add("POST", "/start", lambda q: self.start_task(TrainTask.from_dict(q.json())))
        r.add("POST", "/update/{jobId}", lambda q: self.update_task(TrainTask.from_dict(q.json())))
        r.add("POST", "/metrics/{jobId}",
              lambda q: self.update_metrics(q.params["jobId"], MetricUpdate.from_dict(q.json())))
        r.add("POST", "/finish/{jobId}", self._http_finish)
        r.add("DELETE", "/stop/{jobId}", lambda q: self.stop_task(q.params["jobId"]))
        r.add("GET", "/tasks", lambda q: [t.to_dict() for t in self.list_tasks()])
        r.add("GET", "/metrics", lambda q: (200, self.metrics.render(), "text/plain; version=0.0.4"))
        return r

    def _http_finish(self, q: Request):
        body = q.json() or {}
        err = KubeMLException(body["error"], int(body.get("code", 500))) if body.get("error") else None
        self.job_finished(q.params["jobId"], err)

    def serve(self, host: str, port: int) -> Server:
        srv = Server(self.router(), host, port).start()
        self.url = f"http://{'127.0.0.1' if host in ('0.0.0.0', '') else host}:{srv.port}"
        return srv
