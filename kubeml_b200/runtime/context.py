"""Invocation context: what `flask.request` / `current_app` were to a reference function
(python/kubeml/kubeml/dataset.py:63-78, network.py:140-144).  A worker sets it around each call
of the user's `main()`, so `KubeModel.start()` and `_KubeArgs.parse()` keep their no-argument
signatures."""
from __future__ import annotations

import contextlib
import logging
import threading
from dataclasses import dataclass, field
from typing import Any, Dict, Optional

_tls = threading.local()


@dataclass
class Invocation:
    query: Dict[str, Any]
    body: Any = None                       # JSON body of an infer request
    services: Any = None                   # kubeml_b200.runtime.services.Services
    logger: logging.Logger = field(default_factory=lambda: logging.getLogger("kubeml.function"))
    worker_state: Optional[dict] = None    # warm per-process cache (arena, graphs, pinned data)


def current_invocation() -> Invocation:
    inv = getattr(_tls, "inv", None)
    if inv is None:
        raise RuntimeError("no active KubeML invocation: call through a worker, "
                           "or wrap the call in runtime.context.invocation(...)")
    return inv


def has_invocation() -> bool:
    return getattr(_tls, "inv", None) is not None


@contextlib.contextmanager
def invocation(inv: Invocation):
    prev = getattr(_tls, "inv", None)
    _tls.inv = inv
    try:
        yield inv
    finally:
        _tls.inv = prev
