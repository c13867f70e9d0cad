"""Worker: executes one function invocation (the role of the Fission environment server,
ml/environment/server.py:28-167, minus Flask).

`Worker.invoke()` loads ("specialises") the user function, installs the invocation context,
calls `main()` and maps the outcome onto the reference's wire contract: a successful call yields
`(payload, 200)`; a `KubeMLException` yields its `{error, code}` with its status; anything else a
500 with `repr(e)` (server.py:133-151).
"""
from __future__ import annotations

import logging
import os
import traceback
from typing import Any, Dict, Optional, Tuple

from ..exceptions import KubeMLException, to_error_payload
from .context import Invocation, invocation
from .functions import FunctionRegistry
from .services import Services


class Worker:
    def __init__(self, services: Services, registry: Optional[FunctionRegistry] = None,
                 gpu_id: Optional[int] = None, name: str = "worker"):
        self.services = services
        self.registry = registry or FunctionRegistry()
        self.gpu_id = gpu_id
        self.log = logging.getLogger(f"kubeml.{name}")
        self.state: dict = {}       # warm state reused across invocations (graphs, pinned data)

    def invoke(self, function_name: str, query: Dict[str, Any], body: Any = None
               ) -> Tuple[Dict[str, Any] | list, int]:
        if self.gpu_id is not None:
            self.state["gpu_id"] = self.gpu_id
        inv = Invocation(query=dict(query), body=body, services=self.services,
                         logger=self.log, worker_state=self.state)
        try:
            main = self.registry.load(function_name)
            with invocation(inv):
                res = main()
            if isinstance(res, tuple) and len(res) == 2:
                return res[0], int(res[1])
            return res, 200
        except KubeMLException as e:
            self.log.error("function %s failed: %s", function_name, e.message)
            return e.to_dict(), e.status_code
        except Exception as e:  # noqa: BLE001
            self.log.error("function %s crashed: %s\n%s", function_name, e, traceback.format_exc())
            return to_error_payload(e), 500
