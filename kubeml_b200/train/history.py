"""Persistent training histories: the `kubeml.history` Mongo collection of the reference
(`{_id, task, data}`, ml/pkg/train/util.go:247-280; CRUD in controller/historyApi.go:14-111)
kept as one JSON file per task."""
from __future__ import annotations

import json
import os
import threading
from typing import List, Optional

from ..api.const import home
from ..api.types import History
from ..exceptions import KubeMLException


class HistoryStore:
    def __init__(self, root: Optional[str] = None):
        self.root = root or os.path.join(home(), "history")
        os.makedirs(self.root, exist_ok=True)
        self._lock = threading.Lock()

    def _path(self, task_id: str) -> str:
        if not task_id or "/" in task_id or task_id.startswith("."):
            raise KubeMLException(f"invalid task id {task_id!r}", 400)
        return os.path.join(self.root, f"{task_id}.json")

    def save(self, h: History) -> None:
        with self._lock:
            tmp = self._path(h.id) + ".tmp"
            with open(tmp, "w") as f:
                json.dump(h.to_dict(), f)
            os.replace(tmp, self._path(h.id))

    def get(self, task_id: str) -> History:
        try:
            with open(self._path(task_id)) as f:
                return History.from_dict(json.load(f))
        except FileNotFoundError:
            raise KubeMLException(f"history {task_id} not found", 404)

    def list(self) -> List[History]:
        out = []
        for fn in sorted(os.listdir(self.root)):
            if fn.endswith(".json"):
                try:
                    out.append(self.get(fn[:-5]))
                except (KubeMLException, ValueError):
                    continue
        return out

    def delete(self, task_id: str) -> None:
        try:
            os.remove(self._path(task_id))
        except FileNotFoundError:
            raise KubeMLException(f"history {task_id} not found", 404)

    def prune(self) -> int:
        n = 0
        for fn in os.listdir(self.root):
            if fn.endswith(".json"):
                os.remove(os.path.join(self.root, fn))
                n += 1
        return n
