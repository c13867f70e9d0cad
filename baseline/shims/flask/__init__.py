"""Minimal stand-in for Flask's request context: `request.args.get(name, type=...)`, `request.json`,
`jsonify`, `current_app.logger` — the four things the reference library touches."""
import logging
import threading


class _Args(dict):
    def get(self, key, default=None, type=None):
        v = super().get(key, default)
        if v is None or type is None:
            return v
        return type(v)


class _Request(threading.local):
    def __init__(self):
        self.args = _Args()
        self.json = None


request = _Request()


def set_request(args, json=None):
    request.args = _Args({k: str(v) for k, v in args.items()})
    request.json = json


def jsonify(*a, **k):
    if a and not k:
        return a[0]
    return dict(k)


class _App:
    logger = logging.getLogger("kubeml-reference-function")


current_app = _App()


class Response:  # only used in a type annotation by the reference
    pass
