"""User-visible exception hierarchy (reference: python/kubeml/kubeml/exceptions.py:1-48) and the
`{"error": ..., "code": ...}` JSON contract every service answers with
(ml/pkg/error/error.go:13-87, ml/environment/server.py:133-151)."""
from __future__ import annotations

import json
from typing import Any, Dict, Optional


class KubeMLException(Exception):
    def __init__(self, message: str, status_code: int = 500):
        super().__init__(message)
        self.message = message
        self.status_code = int(status_code)

    def to_dict(self) -> Dict[str, Any]:
        return {"error": self.message, "code": self.status_code}

    def to_json(self) -> str:
        return json.dumps(self.to_dict())

    def __str__(self) -> str:  # error.go:22 Error() returns the message only
        return self.message

    def __reduce__(self):
        # Exceptions cross process boundaries (worker -> control plane); keep both fields.
        return (_rebuild, (self.__class__.__name__, self.message, self.status_code))


def _rebuild(cls_name: str, message: str, code: int) -> "KubeMLException":
    e = KubeMLException(message, code)
    klass = globals().get(cls_name)
    if isinstance(klass, type) and issubclass(klass, KubeMLException):
        e.__class__ = klass
    return e


class MergeError(KubeMLException):
    def __init__(self, e: Optional[BaseException] = None):
        super().__init__(f"Error merging model: {e}", 500)


class DataError(KubeMLException):
    def __init__(self):
        super().__init__("Data not present in request", 400)


class InvalidFormatError(KubeMLException):
    def __init__(self):
        super().__init__("The data provided is not in an appropriate format", 400)


class StorageError(KubeMLException):
    def __init__(self, e: BaseException):
        super().__init__(f"Could not access storage service: {e}", 500)


class DatasetNotFoundError(KubeMLException):
    def __init__(self):
        super().__init__("Dataset not found in storage service", 404)


class InvalidArgsError(KubeMLException):
    def __init__(self, e: BaseException):
        super().__init__(f"Error parsing function arguments: {e}", 500)


class WorkerFault(KubeMLException):
    """A worker process died or missed its round deadline (no reference counterpart: Fission
    reported this as an HTTP 5xx from the router)."""

    def __init__(self, what: str):
        super().__init__(f"Worker fault: {what}", 503)


def error_from_response(status: int, body: bytes | str) -> Optional[KubeMLException]:
    """Decode a service/function reply the way error.go:36-59 does: 200 is success, otherwise
    try `{"code","error"}` JSON and fall back to the trimmed body text."""
    if status == 200:
        return None
    text = body.decode("utf-8", "replace") if isinstance(body, (bytes, bytearray)) else str(body)
    try:
        d = json.loads(text)
        if isinstance(d, dict) and "error" in d:
            return KubeMLException(str(d["error"]), int(d.get("code", status)))
    except (ValueError, TypeError):
        pass
    return KubeMLException(text.strip(), status)


def to_error_payload(exc: BaseException) -> Dict[str, Any]:
    """server.py:133-151: KubeMLException keeps its status, anything else is a 500 with repr()."""
    if isinstance(exc, KubeMLException):
        return exc.to_dict()
    return {"error": repr(exc), "code": 500}
