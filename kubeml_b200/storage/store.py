"""Local shard store.

The reference keeps every dataset as a Mongo database with two collections, `train` and `test`,
one document per 64-sample subset: `{_id: i, data: pickle(x[64i:64i+64]), labels: pickle(...)}`
(python/storage/api.py:105-142, python/storage/utils.py:6-25) and workers fetch a contiguous id
range and re-stack it (python/kubeml/kubeml/dataset.py:184-223).

Here a dataset is a directory holding one contiguous `.npy` per split that is memory-mapped, so
"documents [a, b)" is the zero-copy row slice [64a, 64b).  Subset ids, counts and the contiguous
range semantics are identical; pickling, TCP and `np.vstack` are gone.  For GPU workers the
split can be staged once into page-locked host memory (`pinned()`), from which per-round
slices are copied with `cudaMemcpyAsync` on a side stream.
"""
from __future__ import annotations

import json
import os
import pickle
import shutil
import threading
from typing import Dict, List, Optional, Tuple

import numpy as np

from ..api.const import STORAGE_SUBSET_SIZE, home
from ..api.types import DatasetSummary
from ..exceptions import DatasetNotFoundError, KubeMLException

_SPLITS = ("train", "test")


class ShardStore:
    def __init__(self, root: Optional[str] = None):
        self.root = root or os.path.join(home(), "datasets")
        os.makedirs(self.root, exist_ok=True)
        self._lock = threading.RLock()
        self._mmaps: Dict[Tuple[str, str], Tuple[np.ndarray, np.ndarray]] = {}
        self._pinned: Dict[Tuple[str, str], tuple] = {}
        self._meta: Dict[str, tuple] = {}
        self._resident: Dict[Tuple[str, str, str], tuple] = {}

    # ---- catalogue -------------------------------------------------------------------
    def _dir(self, name: str) -> str:
        if not name or "/" in name or name.startswith("."):
            raise KubeMLException(f"Invalid dataset name {name!r}", 400)
        return os.path.join(self.root, name)

    def exists(self, name: str) -> bool:
        return os.path.isfile(os.path.join(self._dir(name), "meta.json"))

    def list(self) -> List[str]:
        return sorted(d for d in os.listdir(self.root)
                      if os.path.isfile(os.path.join(self.root, d, "meta.json")))

    def meta(self, name: str) -> dict:
        # asked for on every round of every worker: parsed once per version of the file (one stat per call)
        path = os.path.join(self._dir(name), "meta.json")
        try:
            st = os.stat(path)
        except OSError:
            raise DatasetNotFoundError()
        stamp = (st.st_mtime_ns, st.st_size, st.st_ino)
        hit = self._meta.get(name)
        if hit is None or hit[0] != stamp:
            with open(path) as f:
                hit = (stamp, json.load(f))
            self._meta[name] = hit
        return hit[1]

    # ---- ingest ----------------------------------------------------------------------
    def create(self, name: str, x_train, y_train, x_test, y_test,
               subset_size: int = STORAGE_SUBSET_SIZE) -> DatasetSummary:
        """Store four arrays as a dataset.  Duplicate names are a 400 as in api.py:72-74."""
        with self._lock:
            if self.exists(name):
                raise KubeMLException(f"Dataset {name} already exists", 400)
            d = self._dir(name)
            tmp = d + ".partial"
            shutil.rmtree(tmp, ignore_errors=True)
            os.makedirs(tmp)
            meta = {"name": name, "subset_size": int(subset_size)}
            for split, (x, y) in zip(_SPLITS, ((x_train, y_train), (x_test, y_test))):
                x = np.ascontiguousarray(np.asarray(x))
                y = np.ascontiguousarray(np.asarray(y))
                if len(x) != len(y):
                    shutil.rmtree(tmp, ignore_errors=True)
                    raise KubeMLException(f"{split}: features and labels differ in length", 400)
                # labels are flattened on read (dataset.py:213-221): store (n,1) as (n,)
                if y.ndim > 1 and int(np.prod(y.shape[1:])) == 1:
                    y = y.reshape(-1)
                np.save(os.path.join(tmp, f"x-{split}.npy"), x)
                np.save(os.path.join(tmp, f"y-{split}.npy"), y)
                meta[split] = {"samples": int(len(x)),
                               "docs": int(-(-len(x) // subset_size)),
                               "shape": list(x.shape[1:]), "dtype": str(x.dtype),
                               "label_dtype": str(y.dtype)}
            with open(os.path.join(tmp, "meta.json"), "w") as f:
                json.dump(meta, f)
            os.replace(tmp, d)
            return self.summary(name)

    def create_from_files(self, name: str, files: Dict[str, str]) -> DatasetSummary:
        """`files` maps x-train/y-train/x-test/y-test to paths ending in .npy or .pkl
        (api.py:105-124; any other extension is a 400)."""
        arrays = {}
        for key in ("x-train", "y-train", "x-test", "y-test"):
            if key not in files:
                raise KubeMLException("Request does not include a file", 400)
            arrays[key] = load_array_file(files[key])
        return self.create(name, arrays["x-train"], arrays["y-train"],
                           arrays["x-test"], arrays["y-test"])

    def delete(self, name: str) -> None:
        with self._lock:
            if not self.exists(name):
                raise KubeMLException("Dataset does not exist", 404)
            for k in [k for k in self._mmaps if k[0] == name]:
                del self._mmaps[k]
            for k in [k for k in self._pinned if k[0] == name]:
                del self._pinned[k]
            for k in [k for k in self._resident if k[0] == name]:
                del self._resident[k]
            shutil.rmtree(self._dir(name))

    # ---- read ------------------------------------------------------------------------
    def summary(self, name: str) -> DatasetSummary:
        """controller/storageApi.go:69-131 reports docs*64; we report the exact sample count
        (the reference over-counts a short last subset)."""
        m = self.meta(name)
        return DatasetSummary(name=name, train_set_size=m["train"]["samples"],
                              test_set_size=m["test"]["samples"])

    def count_docs(self, name: str, split: str) -> int:
        return int(self.meta(name)[split]["docs"])

    def arrays(self, name: str, split: str) -> Tuple[np.ndarray, np.ndarray]:
        key = (name, split)
        with self._lock:
            if key not in self._mmaps:
                if not self.exists(name):
                    raise DatasetNotFoundError()
                d = self._dir(name)
                self._mmaps[key] = (np.load(os.path.join(d, f"x-{split}.npy"), mmap_mode="r"),
                                    np.load(os.path.join(d, f"y-{split}.npy"), mmap_mode="r"))
            return self._mmaps[key]

    def load_docs(self, name: str, split: str, start: int, end: int
                  ) -> Tuple[np.ndarray, np.ndarray]:
        """Subsets [start, end) in storage order, labels flattened (dataset.py:196-223)."""
        x, y = self.arrays(name, split)
        s = self.meta(name)["subset_size"]
        lo, hi = start * s, min(end * s, len(x))
        return x[lo:hi], np.asarray(y[lo:hi]).reshape(-1) if y.ndim == 1 else np.asarray(y[lo:hi])

    def pinned(self, name: str, split: str):
        """(features, labels) as page-locked torch tensors, staged once per process."""
        import torch
        key = (name, split)
        with self._lock:
            if key not in self._pinned:
                x, y = self.arrays(name, split)
                xt = torch.from_numpy(np.array(x, copy=True))
                yt = torch.from_numpy(np.array(y, copy=True))
                if torch.cuda.is_available():
                    xt, yt = xt.pin_memory(), yt.pin_memory()
                self._pinned[key] = (xt, yt)
            return self._pinned[key]


    def resident(self, name: str, split: str, device):
        """(features, labels) of a whole split as tensors on `device`, staged once per process: for kernel-only
        measurements that want no host->device traffic inside the timed region (`device_resident_data`)."""
        key = (name, split, str(device))
        with self._lock:
            if key not in self._resident:
                px, py = self.pinned(name, split)
                self._resident[key] = (px.to(device), py.to(device))
            return self._resident[key]


def load_array_file(path: str) -> np.ndarray:
    ext = path.rsplit(".", 1)[-1].lower()
    if ext == "npy":
        return np.load(path, allow_pickle=False)
    if ext == "pkl":
        with open(path, "rb") as f:
            return np.asarray(pickle.load(f))
    raise KubeMLException("File extension not supported, must be one of [npy, pkl]", 400)


_default: Optional[ShardStore] = None
_default_lock = threading.Lock()


def default_store() -> ShardStore:
    global _default
    with _default_lock:
        if _default is None or _default.root != os.path.join(home(), "datasets"):
            _default = ShardStore()
        return _default


def set_default_store(store: Optional[ShardStore]) -> None:
    global _default
    with _default_lock:
        _default = store
