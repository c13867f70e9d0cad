"""Minimal JSON-over-HTTP plumbing shared by every role.

The reference wires its services with gorilla/mux (controller/api.go:16-42, scheduler/api.go:184-192,
ps/api.go:335-345, train/api.go:141-149) and Flask (python/storage/api.py, ml/environment/server.py).
On a single host those hops are optional: every role is a Python object that can be called
in-process, and `serve()` exposes the same routes over the stdlib HTTP server when a remote
client (the CLI on another shell, Prometheus, curl) wants them.
"""
from __future__ import annotations

import json
import re
import threading
import urllib.error
import urllib.request
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Any, Callable, Dict, List, Optional, Tuple
from urllib.parse import parse_qs, urlparse

from .exceptions import KubeMLException, error_from_response, to_error_payload


class Request:
    def __init__(self, method: str, path: str, query: Dict[str, str], body: bytes,
                 headers: Dict[str, str], params: Dict[str, str]):
        self.method, self.path, self.query = method, path, query
        self.body, self.headers, self.params = body, headers, params

    def json(self) -> Any:
        if not self.body:
            return None
        return json.loads(self.body.decode("utf-8"))


Handler = Callable[[Request], Any]


class Router:
    """`add("POST", "/dataset/{name}", fn)`; fn returns a JSON-able object, bytes/str, or a
    (status, payload[, content_type]) tuple."""

    def __init__(self):
        self._routes: List[Tuple[str, "re.Pattern[str]", Handler]] = []

    def add(self, method: str, pattern: str, fn: Handler) -> None:
        rx = re.compile("^" + re.sub(r"\{(\w+)\}", r"(?P<\1>[^/]+)", pattern.rstrip("/")) + "/?$")
        self._routes.append((method.upper(), rx, fn))

    def dispatch(self, method: str, url: str, body: bytes = b"",
                 headers: Optional[Dict[str, str]] = None) -> Tuple[int, bytes, str]:
        u = urlparse(url)
        path = u.path or "/"
        query = {k: v[-1] for k, v in parse_qs(u.query).items()}
        seen_path = False
        for m, rx, fn in self._routes:
            mt = rx.match(path)
            if not mt:
                continue
            seen_path = True
            if m != method.upper():
                continue
            req = Request(method.upper(), path, query, body, headers or {}, mt.groupdict())
            try:
                return _encode(fn(req))
            except Exception as e:  # noqa: BLE001 - the contract maps *every* failure to JSON
                p = to_error_payload(e)
                return int(p["code"]), json.dumps(p).encode(), "application/json"
        code = 405 if seen_path else 404
        return code, json.dumps({"error": "not found" if code == 404 else "method not allowed",
                                 "code": code}).encode(), "application/json"


def _encode(res: Any) -> Tuple[int, bytes, str]:
    status, ctype = 200, "application/json"
    if isinstance(res, tuple):
        if len(res) == 3:
            status, res, ctype = res
        else:
            status, res = res
    if isinstance(res, (bytes, bytearray)):
        return status, bytes(res), ctype if ctype != "application/json" else "application/octet-stream"
    if isinstance(res, str) and ctype != "application/json":
        return status, res.encode(), ctype
    if res is None:
        return status, b"", ctype
    if hasattr(res, "to_dict"):
        res = res.to_dict()
    return status, json.dumps(res).encode(), ctype


class Server:
    def __init__(self, router: Router, host: str, port: int):
        outer = router

        class _H(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"

            def _go(self):
                n = int(self.headers.get("Content-Length") or 0)
                body = self.rfile.read(n) if n else b""
                status, payload, ctype = outer.dispatch(self.command, self.path, body,
                                                        dict(self.headers.items()))
                self.send_response(status)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(payload)))
                self.end_headers()
                self.wfile.write(payload)

            do_GET = do_POST = do_DELETE = do_PUT = _go

            def log_message(self, *a):  # quiet
                pass

        self.httpd = ThreadingHTTPServer((host, port), _H)
        self.httpd.daemon_threads = True
        self.port = self.httpd.server_address[1]
        self._thread: Optional[threading.Thread] = None

    def start(self) -> "Server":
        self._thread = threading.Thread(target=self.httpd.serve_forever, daemon=True,
                                        name=f"httpd:{self.port}")
        self._thread.start()
        return self

    def stop(self) -> None:
        self.httpd.shutdown()
        self.httpd.server_close()


def http_call(method: str, url: str, payload: Any = None, raw: Optional[bytes] = None,
              headers: Optional[Dict[str, str]] = None, timeout: float = 3600.0) -> Any:
    """Client side of the contract: JSON in, JSON out, non-200 raised as KubeMLException
    (ml/pkg/error/error.go:36-59)."""
    data = raw
    hdrs = dict(headers or {})
    if payload is not None:
        if hasattr(payload, "to_dict"):
            payload = payload.to_dict()
        data = json.dumps(payload).encode()
        hdrs.setdefault("Content-Type", "application/json")
    req = urllib.request.Request(url, data=data, method=method.upper(), headers=hdrs)
    try:
        with urllib.request.urlopen(req, timeout=timeout) as r:
            status, body, ctype = r.status, r.read(), r.headers.get("Content-Type", "")
    except urllib.error.HTTPError as e:
        status, body, ctype = e.code, e.read(), e.headers.get("Content-Type", "")
    except (urllib.error.URLError, OSError) as e:
        raise KubeMLException(f"could not reach {url}: {e}", 503)
    err = error_from_response(status, body)
    if err is not None:
        raise err
    if not body:
        return None
    if "json" in ctype:
        return json.loads(body.decode())
    return body


def encode_multipart(files: Dict[str, Tuple[str, bytes]]) -> Tuple[bytes, str]:
    """multipart/form-data body for dataset upload (controller/client/v1/dataset.go:50-106)."""
    boundary = "kubemlb200boundary7d3f"
    out = bytearray()
    for field, (filename, content) in files.items():
        out += f"--{boundary}\r\n".encode()
        out += (f'Content-Disposition: form-data; name="{field}"; filename="{filename}"\r\n'
                "Content-Type: application/octet-stream\r\n\r\n").encode()
        out += content + b"\r\n"
    out += f"--{boundary}--\r\n".encode()
    return bytes(out), f"multipart/form-data; boundary={boundary}"


def decode_multipart(body: bytes, content_type: str) -> Dict[str, Tuple[str, bytes]]:
    m = re.search(r"boundary=([^;]+)", content_type or "")
    if not m:
        return {}
    boundary = ("--" + m.group(1).strip('"')).encode()
    out: Dict[str, Tuple[str, bytes]] = {}
    for part in body.split(boundary):
        part = part.strip(b"\r\n")
        if not part or part == b"--":
            continue
        head, _, content = part.partition(b"\r\n\r\n")
        hm = re.search(rb'name="([^"]+)"(?:; filename="([^"]*)")?', head)
        if not hm:
            continue
        out[hm.group(1).decode()] = ((hm.group(2) or b"").decode(), content)
    return out
