"""Checkpoint save / load (SURVEY 5.4, W24 — the reference has no persistent model: its final
`clearTensors` deletes the reference model with the replicas, ml/pkg/train/util.go:208-244).

A checkpoint is a directory `<root>/<jobId>/` with
  * `model.pt`     — `torch.save` of the reference `state_dict` (float32 tensors, int64 BatchNorm
                     counters), loadable with `nn.Module.load_state_dict`;
  * `index.json`   — the `{jobId}:{name}` key layout: name, dtype, shape per entry in state_dict order;
  * `history.json` — the task's `History` document.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np
import torch

from ..api.types import History, JobHistory, TrainRequest
from ..model.store import TensorStore, weight_key


def save_state_dict(root: str, job_id: str, state: Dict[str, torch.Tensor],
                    request: Optional[TrainRequest] = None,
                    history: Optional[JobHistory] = None) -> str:
    d = os.path.join(root, job_id)
    os.makedirs(d, exist_ok=True)
    cpu = {k: v.detach().to("cpu").contiguous() for k, v in state.items()}
    torch.save(cpu, os.path.join(d, "model.pt"))
    with open(os.path.join(d, "index.json"), "w") as f:
        json.dump([{"key": weight_key(job_id, k), "name": k, "dtype": str(v.dtype).replace("torch.", ""),
                    "shape": list(v.shape)} for k, v in cpu.items()], f)
    if request is not None or history is not None:
        with open(os.path.join(d, "history.json"), "w") as f:
            json.dump(History(id=job_id, task=request or TrainRequest(),
                              data=history or JobHistory()).to_dict(), f)
    return d


def save_checkpoint(root: str, job_id: str, store: TensorStore, layer_names: Iterable[str],
                    request: Optional[TrainRequest] = None,
                    history: Optional[JobHistory] = None) -> str:
    state = {n: torch.from_numpy(np.array(store.tensorget(weight_key(job_id, n)), copy=True))
             for n in layer_names}
    return save_state_dict(root, job_id, state, request, history)


def load_checkpoint(root: str, job_id: str) -> Tuple[Dict[str, torch.Tensor], Optional[History]]:
    d = os.path.join(root, job_id)
    state = torch.load(os.path.join(d, "model.pt"), map_location="cpu", weights_only=True)
    hist = None
    hp = os.path.join(d, "history.json")
    if os.path.isfile(hp):
        with open(hp) as f:
            hist = History.from_dict(json.load(f))
    return state, hist


def restore_into_store(root: str, job_id: str, store: TensorStore,
                       as_job: Optional[str] = None) -> List[str]:
    """Publish a checkpoint as the reference model of `as_job` (resume / warm start)."""
    state, _ = load_checkpoint(root, job_id)
    target = as_job or job_id
    store.multi_set({weight_key(target, n): t.numpy() for n, t in state.items()})
    return list(state.keys())
