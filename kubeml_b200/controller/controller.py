"""Controller: the stateless front door (reference: ml/pkg/controller/api.go:16-42 and
networkApi.go, storageApi.go, historyApi.go, tasksApi.go).

Routes kept verbatim:
    POST /train, POST /infer,
    GET|POST|DELETE /dataset/{name}, GET /dataset,
    GET /tasks, DELETE /tasks/{jobId},
    GET|DELETE /history/{taskId}, GET|DELETE /history,
    GET /health
plus `GET|POST /function`, `DELETE /function/{name}` (the reference CLI talks to the Fission
API for these, cmd/function.go) and `GET /metrics` for convenience.
"""
from __future__ import annotations

import logging
from typing import Any, Dict, List, Optional, Tuple

from ..api.types import DatasetSummary, History, InferRequest, TrainRequest, TrainTask
from ..exceptions import KubeMLException
from ..httpd import Request, Router, Server, decode_multipart
from ..runtime.functions import FunctionRegistry
from ..storage.service import StorageService
from ..train.history import HistoryStore


class Controller:
    def __init__(self, scheduler: Any, ps: Any, storage: StorageService, histories: HistoryStore,
                 functions: Optional[FunctionRegistry] = None):
        self.scheduler, self.ps, self.storage = scheduler, ps, storage
        self.histories = histories
        self.functions = functions or FunctionRegistry()
        self.log = logging.getLogger("kubeml.controller")

    # ---- networks (networkApi.go:12-72) --------------------------------------------------------
    def train(self, req: TrainRequest) -> str:
        return self.scheduler.submit_train_task(req)

    def infer(self, req: InferRequest) -> Any:
        return self.scheduler.infer(req)

    # ---- datasets (storageApi.go:35-189) -------------------------------------------------------
    def create_dataset(self, name: str, files: Dict[str, Tuple[str, bytes]]) -> dict:
        return self.storage.upload(name, files)

    def delete_dataset(self, name: str) -> dict:
        return self.storage.delete(name)

    def get_dataset(self, name: str) -> DatasetSummary:
        return self.storage.store.summary(name)

    def list_datasets(self) -> List[DatasetSummary]:
        return [self.storage.store.summary(n) for n in self.storage.store.list()]

    # ---- histories (historyApi.go:14-111) ------------------------------------------------------
    def get_history(self, task_id: str) -> History:
        return self.histories.get(task_id)

    def list_histories(self) -> List[History]:
        return self.histories.list()

    def delete_history(self, task_id: str) -> None:
        self.histories.delete(task_id)

    def prune_histories(self) -> int:
        return self.histories.prune()

    # ---- tasks (tasksApi.go:10-36) -------------------------------------------------------------
    def list_tasks(self) -> List[TrainTask]:
        return self.ps.list_tasks()

    def stop_task(self, job_id: str) -> None:
        self.ps.stop_task(job_id)

    # ---- HTTP ----------------------------------------------------------------------------------
    def router(self) -> Router:
        r = Router()
        r.add("GET", "/health", lambda q: (200, None))
        r.add("POST", "/train", lambda q: (200, self.train(TrainRequest.from_dict(q.json())), "text/plain"))
        r.add("POST", "/infer", lambda q: self.infer(InferRequest.from_dict(q.json())))
        r.add("GET", "/dataset", lambda q: [d.to_dict() for d in self.list_datasets()])
        r.add("GET", "/dataset/{name}", lambda q: self.get_dataset(q.params["name"]))
        r.add("POST", "/dataset/{name}", self._http_create_dataset)
        r.add("DELETE", "/dataset/{name}", lambda q: self.delete_dataset(q.params["name"]))
        r.add("GET", "/tasks", lambda q: [t.to_dict() for t in self.list_tasks()])
        r.add("DELETE", "/tasks/{jobId}", lambda q: self.stop_task(q.params["jobId"]))
        r.add("GET", "/history", lambda q: [h.to_dict() for h in self.list_histories()])
        r.add("DELETE", "/history", lambda q: {"deleted": self.prune_histories()})
        r.add("GET", "/history/{taskId}", lambda q: self.get_history(q.params["taskId"]))
        r.add("DELETE", "/history/{taskId}", lambda q: self.delete_history(q.params["taskId"]))
        r.add("GET", "/function", lambda q: {"functions": self.functions.list(),
                                             "builtin": self.functions.builtins()})
        r.add("POST", "/function/{name}", self._http_create_function)
        r.add("DELETE", "/function/{name}", lambda q: self.functions.delete(q.params["name"]))
        r.add("GET", "/metrics", lambda q: (200, self.ps.metrics.render(), "text/plain; version=0.0.4"))
        return r

    def _http_create_dataset(self, q: Request):
        ctype = q.headers.get("Content-Type") or q.headers.get("content-type") or ""
        return self.create_dataset(q.params["name"], decode_multipart(q.body, ctype))

    def _http_create_function(self, q: Request):
        import os
        import tempfile
        body = q.json() or {}
        code = body.get("code")
        if not code:
            raise KubeMLException("function code missing", 400)
        with tempfile.NamedTemporaryFile("w", suffix=".py", delete=False) as f:
            f.write(code)
        try:
            self.functions.create(q.params["name"], f.name)
        finally:
            os.remove(f.name)

    def serve(self, host: str, port: int) -> Server:
        return Server(self.router(), host, port).start()
