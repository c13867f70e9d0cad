"""Function environment server: the Fission "environment" contract of the reference over HTTP
(ml/environment/server.py:28-167) — `/specialize` and `/v2/specialize` load a user function,
`/healthz` answers 200, and `/` runs the loaded function with the request's query string as its
arguments.  The worker processes do not need HTTP to be driven (they are fed by the train job),
this server exists so that a function can also be exercised with curl exactly like a Fission pod."""
from __future__ import annotations

import os
from typing import Optional

from ..exceptions import KubeMLException
from ..httpd import Request, Router, Server
from .functions import FunctionRegistry
from .services import Services
from .worker import Worker


class FunctionServer:
    def __init__(self, services: Services, registry: Optional[FunctionRegistry] = None, gpu_id: Optional[int] = None):
        self.worker = Worker(services, registry, gpu_id=gpu_id, name="funcserver")
        self.function: Optional[str] = None

    def specialize(self, req: Request):
        """v1: the function lives at /userfunc/user; here the body names a registered function."""
        body = req.json() or {}
        name = body.get("functionName") or body.get("filepath") or os.environ.get("KUBEML_FUNCTION")
        if not name:
            raise KubeMLException("specialize needs functionName", 400)
        if os.path.isfile(name):                      # v2: {"filepath": ..., "functionName": "module.func"}
            reg = self.worker.registry
            fname = os.path.splitext(os.path.basename(name))[0]
            if not reg.exists(fname):
                reg.create(fname, name)
            name = fname
        if not self.worker.registry.exists(name):
            raise KubeMLException(f"function {name} does not exist", 404)
        self.function = name
        return {"specialized": name}

    def call(self, req: Request):
        if self.function is None:
            raise KubeMLException("Generic container: no requests supported", 500)
        payload, status = self.worker.invoke(self.function, req.query, req.json() if req.body else None)
        return status, payload

    def router(self) -> Router:
        r = Router()
        r.add("POST", "/specialize", self.specialize)
        r.add("POST", "/v2/specialize", self.specialize)
        r.add("GET", "/healthz", lambda q: (200, None))
        for m in ("GET", "POST", "PUT", "DELETE"):
            r.add(m, "/", self.call)
        return r

    def serve(self, host: str = "127.0.0.1", port: int = 8888) -> Server:
        return Server(self.router(), host, port).start()
