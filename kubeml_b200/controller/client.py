"""Typed client SDK (reference: ml/pkg/controller/client/client.go:31-59 and v1/*.go):

    c = make_kubeml_client()            # HTTP to a running control plane, or
    c = make_kubeml_client(cluster)     # direct calls into an in-process Cluster
    c.v1().networks().train(req) / .infer(req)
    c.v1().datasets().create(name, x_train, y_train, x_test, y_test) / .delete / .get / .list
    c.v1().histories().get(id) / .delete(id) / .list() / .prune()
    c.v1().tasks().list() / .stop(id)
    c.v1().functions().create(name, path) / .delete(name) / .list()

The reference discovers the controller through kubeconfig + the `controller` Service
(client/util.go:17-65); here it is `KUBEML_CONTROLLER_URL` or http://127.0.0.1:10100.
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Optional, Tuple

from ..api import const
from ..api.types import DatasetSummary, History, InferRequest, TrainRequest, TrainTask
from ..httpd import encode_multipart, http_call


class _Transport:
    def __init__(self, cluster=None, url: Optional[str] = None):
        self.cluster = cluster
        self.url = (url or os.environ.get("KUBEML_CONTROLLER_URL") or const.CONTROLLER_URL).rstrip("/")

    @property
    def local(self) -> bool:
        return self.cluster is not None


class Networks:
    def __init__(self, t: _Transport):
        self.t = t

    def train(self, req: TrainRequest) -> str:
        if self.t.local:
            return self.t.cluster.controller.train(req)
        out = http_call("POST", f"{self.t.url}/train", req)
        return out.decode() if isinstance(out, bytes) else str(out)

    def infer(self, req: InferRequest) -> Any:
        if self.t.local:
            return self.t.cluster.controller.infer(req)
        return http_call("POST", f"{self.t.url}/infer", req)


class Datasets:
    def __init__(self, t: _Transport):
        self.t = t

    def create(self, name: str, x_train: str, y_train: str, x_test: str, y_test: str) -> dict:
        """Arguments are file paths (.npy/.pkl), uploaded as the four multipart parts
        x-train/y-train/x-test/y-test (v1/dataset.go:50-106)."""
        files: Dict[str, Tuple[str, bytes]] = {}
        for field, path in (("x-train", x_train), ("y-train", y_train),
                            ("x-test", x_test), ("y-test", y_test)):
            with open(path, "rb") as f:
                files[field] = (os.path.basename(path), f.read())
        if self.t.local:
            return self.t.cluster.controller.create_dataset(name, files)
        body, ctype = encode_multipart(files)
        return http_call("POST", f"{self.t.url}/dataset/{name}", raw=body,
                         headers={"Content-Type": ctype})

    def delete(self, name: str) -> dict:
        if self.t.local:
            return self.t.cluster.controller.delete_dataset(name)
        return http_call("DELETE", f"{self.t.url}/dataset/{name}")

    def get(self, name: str) -> DatasetSummary:
        if self.t.local:
            return self.t.cluster.controller.get_dataset(name)
        return DatasetSummary.from_dict(http_call("GET", f"{self.t.url}/dataset/{name}"))

    def list(self) -> List[DatasetSummary]:
        if self.t.local:
            return self.t.cluster.controller.list_datasets()
        return [DatasetSummary.from_dict(d) for d in http_call("GET", f"{self.t.url}/dataset") or []]


class Histories:
    def __init__(self, t: _Transport):
        self.t = t

    def get(self, task_id: str) -> History:
        if self.t.local:
            return self.t.cluster.controller.get_history(task_id)
        return History.from_dict(http_call("GET", f"{self.t.url}/history/{task_id}"))

    def delete(self, task_id: str) -> None:
        if self.t.local:
            return self.t.cluster.controller.delete_history(task_id)
        http_call("DELETE", f"{self.t.url}/history/{task_id}")

    def list(self) -> List[History]:
        if self.t.local:
            return self.t.cluster.controller.list_histories()
        return [History.from_dict(d) for d in http_call("GET", f"{self.t.url}/history") or []]

    def prune(self) -> None:
        if self.t.local:
            self.t.cluster.controller.prune_histories()
            return
        http_call("DELETE", f"{self.t.url}/history")


class Tasks:
    def __init__(self, t: _Transport):
        self.t = t

    def list(self) -> List[TrainTask]:
        if self.t.local:
            return self.t.cluster.controller.list_tasks()
        return [TrainTask.from_dict(d) for d in http_call("GET", f"{self.t.url}/tasks") or []]

    def stop(self, job_id: str) -> None:
        if self.t.local:
            return self.t.cluster.controller.stop_task(job_id)
        http_call("DELETE", f"{self.t.url}/tasks/{job_id}")


class Functions:
    def __init__(self, t: _Transport):
        self.t = t

    def create(self, name: str, code_path: str) -> None:
        if self.t.local:
            return self.t.cluster.functions.create(name, code_path)
        with open(code_path) as f:
            http_call("POST", f"{self.t.url}/function/{name}", {"code": f.read()})

    def delete(self, name: str) -> None:
        if self.t.local:
            return self.t.cluster.functions.delete(name)
        http_call("DELETE", f"{self.t.url}/function/{name}")

    def list(self) -> dict:
        if self.t.local:
            return {"functions": self.t.cluster.functions.list(),
                    "builtin": self.t.cluster.functions.builtins()}
        return http_call("GET", f"{self.t.url}/function")

    def exists(self, name: str) -> bool:
        if self.t.local:
            return self.t.cluster.functions.exists(name)
        got = self.list()
        short = name.split(":", 1)[-1]
        return (name in got["functions"] or short in got["builtin"]
                or f"function_{short}" in got["builtin"])


class V1:
    def __init__(self, t: _Transport):
        self._t = t

    def networks(self) -> Networks:
        return Networks(self._t)

    def datasets(self) -> Datasets:
        return Datasets(self._t)

    def histories(self) -> Histories:
        return Histories(self._t)

    def tasks(self) -> Tasks:
        return Tasks(self._t)

    def functions(self) -> Functions:
        return Functions(self._t)


class KubemlClient:
    def __init__(self, cluster=None, url: Optional[str] = None):
        self._t = _Transport(cluster, url)

    def v1(self) -> V1:
        return V1(self._t)

    V1 = v1   # the reference spells it client.V1() (client.go:55-59)


def make_kubeml_client(cluster=None, url: Optional[str] = None) -> KubemlClient:
    return KubemlClient(cluster, url)
