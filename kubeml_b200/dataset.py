"""`KubeDataset`: the user's view of a stored dataset (reference:
python/kubeml/kubeml/dataset.py:81-227).

Same contract: subclass it, call `super().__init__("<dataset name>")`, read `self.data` /
`self.labels` (numpy, refilled by the framework with a contiguous range of 64-sample subsets
before every sync round), use `self.is_training()` to pick transforms, and implement
`__getitem__` / `__len__`.

B200-first additions (optional, the reference has no counterpart):

* `transform_batch(x, y)` — if a subclass defines it, batches are cut as contiguous slices of
  the round's data (no per-sample Python, no collate), copied host->device from page-locked
  memory with `cudaMemcpyAsync` on a side stream one batch ahead of compute, and the hook runs
  on the device tensors (normalise / cast / augment there).
"""
from __future__ import annotations

import logging
from abc import ABC, abstractmethod
from typing import Iterator, Optional

import numpy as np
import torch
import torch.utils.data as data

from .args import _KubeArgs  # noqa: F401  (re-exported: reference users import it from here)
from .exceptions import DatasetNotFoundError, StorageError


class KubeDataset(data.Dataset, ABC):
    def __init__(self, dataset: str):
        from .runtime.context import current_invocation, has_invocation
        from .storage import default_store

        self.dataset = dataset
        self._mode: Optional[str] = None
        self._args = None
        self._store = (current_invocation().services.shard_store
                       if has_invocation() and current_invocation().services is not None
                       else default_store())
        self.data, self.labels = None, None
        # page-locked views of the same rows, set when the split was staged for a GPU worker
        self._pinned_data, self._pinned_labels = None, None
        self._want_pinned = False
        self._device_resident = False

        try:
            if not self._store.exists(dataset):
                logging.error("Dataset not in the storage service. Dataset = %s, Available = %s",
                              dataset, self._store.list())
                raise DatasetNotFoundError()
            # dataset.py:121-123: document counts size the epoch
            self.num_docs = self._store.count_docs(dataset, "train")
            self.num_val_docs = self._store.count_docs(dataset, "test")
        except OSError as e:
            raise StorageError(e)

    # ---- mode ----------------------------------------------------------------------------
    def _eval(self) -> None:
        self._mode = "val"

    def _train(self) -> None:
        self._mode = "train"

    def is_training(self) -> bool:
        return self._mode == "train"

    # ---- loading (dataset.py:150-223) ----------------------------------------------------------
    def _load_train_data(self, start: int, end: int) -> None:
        self.data, self.labels = self.__load_data(range(start, end))
        self._train()

    def _load_validation_data(self, start: int, end: int) -> None:
        self.data, self.labels = self.__load_data(range(start, end), validation=True)
        self._eval()

    def _close(self) -> None:
        pass

    def __load_data(self, minibatches: range, validation: bool = False):
        split = "test" if validation else "train"
        try:
            x, y = self._store.load_docs(self.dataset, split, minibatches.start, minibatches.stop)
            if self._want_pinned:
                px, py = self._store.pinned(self.dataset, split)
                s = self._store.meta(self.dataset)["subset_size"]
                lo, hi = minibatches.start * s, min(minibatches.stop * s, len(px))
                self._pinned_data, self._pinned_labels = px[lo:hi], py[lo:hi].reshape(-1)
            else:
                self._pinned_data = self._pinned_labels = None
            self._split, self._rows = split, (minibatches.start, minibatches.stop)
        except OSError as e:
            raise StorageError(e)
        return x, np.asarray(y).reshape(-1)

    @abstractmethod
    def __len__(self):
        ...

    # ---- batching ------------------------------------------------------------------------
    def _has_batch_transform(self) -> bool:
        return callable(getattr(self, "transform_batch", None))

    def _batches(self, batch_size: int, device: torch.device) -> "BatchSource":
        if self._has_batch_transform():
            return SlicedBatches(self, batch_size, device)
        return LoaderBatches(self, batch_size)


class BatchSource:
    def __len__(self) -> int: ...
    def __iter__(self) -> Iterator: ...
    h2d_bytes = 0


class LoaderBatches(BatchSource):
    """Reference behaviour: `DataLoader(dataset, batch_size)` — no shuffle, no workers
    (network.py:284)."""

    def __init__(self, ds: KubeDataset, batch_size: int):
        self.loader = data.DataLoader(ds, batch_size=batch_size)

    def __len__(self):
        return len(self.loader)

    def __iter__(self):
        return iter(self.loader)


class SlicedBatches(BatchSource):
    """Contiguous slices; on CUDA, double-buffered async H2D from pinned memory on a copy
    stream, overlapped with the previous batch's compute."""

    def __init__(self, ds: KubeDataset, batch_size: int, device: torch.device):
        self.ds, self.bs, self.device = ds, int(batch_size), device
        self.n = len(ds.data)
        self.h2d_bytes = 0
        self._resident = False
        self.trace = [] if getattr(ds, "_trace", False) else None     # (batch, start event, end event) of each H2D copy
        if device.type == "cuda" and getattr(ds, "_device_resident", False):
            # kernel-only measurements: the whole split sits on the device (staged once per process), a round's
            # batches are views of it
            rx, ry = ds._store.resident(ds.dataset, ds._split, device)
            sub = ds._store.meta(ds.dataset)["subset_size"]
            lo, hi = ds._rows[0] * sub, min(ds._rows[1] * sub, len(rx))
            self.x, self.y = rx[lo:hi], ry[lo:hi].reshape(-1)
            self._resident = True
        elif device.type == "cuda" and ds._pinned_data is not None:
            self.x, self.y = ds._pinned_data, ds._pinned_labels
        else:
            self.x = torch.from_numpy(np.array(ds.data, copy=True))
            self.y = torch.from_numpy(np.array(ds.labels, copy=True))
        self._copy_stream = None
        self._first = None
        if device.type == "cuda":
            self._copy_stream = _copy_stream_for(device)

    def __len__(self):
        return -(-self.n // self.bs)

    def _stage(self, i: int):
        lo, hi = i * self.bs, min((i + 1) * self.bs, self.n)
        xb, yb = self.x[lo:hi], self.y[lo:hi]
        if self.device.type != "cuda" or self._resident:
            return xb, yb, None
        with torch.cuda.stream(self._copy_stream):
            t0 = None
            if self.trace is not None:         # timed events only when a trace was asked for
                t0 = torch.cuda.Event(enable_timing=True)
                t0.record(self._copy_stream)
            xd = xb.to(self.device, non_blocking=True)
            yd = yb.to(self.device, non_blocking=True)
            ev = torch.cuda.Event(enable_timing=t0 is not None)
            ev.record(self._copy_stream)
            if t0 is not None:
                self.trace.append((i, t0, ev))
        self.h2d_bytes += xb.numel() * xb.element_size() + yb.numel() * yb.element_size()
        return xd, yd, ev

    def prestage(self) -> None:
        """Start the copy of the first batch now (the caller is about to wait for the previous round)."""
        if self._first is None and len(self) > 0:
            self._first = self._stage(0)

    def __iter__(self):
        nb = len(self)
        if nb == 0:
            return
        nxt, self._first = (self._first if self._first is not None else self._stage(0)), None
        for i in range(nb):
            xd, yd, ev = nxt
            if i + 1 < nb:
                nxt = self._stage(i + 1)
            if ev is not None:
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(ev)
                xd.record_stream(cur)
                yd.record_stream(cur)
            yield self.ds.transform_batch(xd, yd)


_copy_streams = {}


def _copy_stream_for(device: torch.device):
    key = (device.type, device.index)
    if key not in _copy_streams:
        _copy_streams[key] = torch.cuda.Stream(device=device)
    return _copy_streams[key]
