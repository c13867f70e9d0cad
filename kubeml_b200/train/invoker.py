"""Function fan-out: how a train job runs N copies of the user function.

Reference: N goroutines each doing an HTTP GET to the Fission router and parsing the JSON reply
(ml/pkg/train/function.go:103-227).  Here an *invoker* owns N workers and runs an invocation on
each concurrently:

* `ThreadInvoker`  — workers are threads of this process (CPU jobs, tests, single-process
                     multi-GPU); torch releases the GIL inside kernels.
* `SpmdInvoker`    — kubeml_b200.runtime.spmd: one OS process per GPU launched by torchrun;
                     rank 0 hosts the job and broadcasts invocations.
* `ProcessInvoker` — kubeml_b200.runtime.pool: the control plane spawns one persistent process
                     per GPU and talks to them over pipes.
"""
from __future__ import annotations

import threading
from typing import Any, Callable, Dict, List, Optional, Tuple

from ..exceptions import KubeMLException, error_from_response
from ..runtime.functions import FunctionRegistry
from ..runtime.services import Services
from ..runtime.worker import Worker

Result = Tuple[int, Optional[dict], Optional[KubeMLException]]   # (func_id, payload, error)


def _payload_error(payload: Any, status: int) -> Optional[KubeMLException]:
    if status == 200:
        return None
    if isinstance(payload, dict) and "error" in payload:
        return KubeMLException(str(payload["error"]), int(payload.get("code", status)))
    return error_from_response(status, str(payload))


class Invoker:
    max_parallelism: int = 1

    def invoke_one(self, func_id: int, function: str, query: Dict[str, Any], body: Any = None
                   ) -> Tuple[Any, int]:
        raise NotImplementedError

    def invoke_many(self, function: str, queries: List[Dict[str, Any]],
                    on_done: Optional[Callable[[int, bool], None]] = None) -> List[Result]:
        raise NotImplementedError

    def close(self) -> None:
        pass


class ThreadInvoker(Invoker):
    def __init__(self, services: Services, registry: Optional[FunctionRegistry] = None,
                 max_parallelism: int = 64, gpu_ids: Optional[List[int]] = None):
        self.services = services
        self.registry = registry or FunctionRegistry()
        self.max_parallelism = max_parallelism
        self.gpu_ids = gpu_ids
        self._workers: Dict[int, Worker] = {}
        if gpu_ids is not None and len(gpu_ids) > 1:
            # several GPU workers as threads of one process: stream capture is per-process state in
            # practice (allocator, generators), so step graphs are left to the process-per-GPU runtime
            self.services.options["cuda_graphs"] = False

    def _worker(self, func_id: int) -> Worker:
        w = self._workers.get(func_id)
        if w is None:
            gpu = self.gpu_ids[func_id % len(self.gpu_ids)] if self.gpu_ids else None
            w = Worker(self.services, self.registry, gpu_id=gpu, name=f"worker{func_id}")
            self._workers[func_id] = w
        return w

    def invoke_one(self, func_id, function, query, body=None):
        return self._worker(func_id).invoke(function, query, body)

    def invoke_many(self, function, queries, on_done=None):
        results: List[Optional[Result]] = [None] * len(queries)

        def run(i: int, q: Dict[str, Any]):
            fid = int(q.get("funcId", i))
            ok = False
            try:
                payload, status = self._worker(fid).invoke(function, q)
                err = _payload_error(payload, status)
                ok = err is None
                results[i] = (fid, payload if ok else None, err)
            except BaseException as e:  # noqa: BLE001 - a worker thread must always report
                results[i] = (fid, None, KubeMLException(repr(e), 500))
            finally:
                if on_done is not None:
                    on_done(fid, ok)

        threads = [threading.Thread(target=run, args=(i, q), name=f"fn-{i}", daemon=True)
                   for i, q in enumerate(queries)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        return [r for r in results if r is not None]
