"""Dataclasses that serialise to exactly the JSON the reference services exchange.

Field names follow the `json:"..."` tags in ml/pkg/api/types.go:13-111 (including the
reference's `validations_loss` spelling in MetricUpdate, types.go:86) so that a client written
against the reference REST API keeps working.
"""
from __future__ import annotations

import dataclasses
import json
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional


def _plain(obj: Any) -> Any:
    if dataclasses.is_dataclass(obj):
        return obj.to_dict()
    if isinstance(obj, (list, tuple)):
        return [_plain(o) for o in obj]
    if isinstance(obj, dict):
        return {k: _plain(v) for k, v in obj.items()}
    return obj


class _Wire:
    """Mixin: JSON names live in field metadata; unknown keys are ignored on input."""

    def to_dict(self) -> Dict[str, Any]:
        out = {}
        for f in dataclasses.fields(self):
            if f.metadata.get("skip"):
                continue
            if f.metadata.get("omitempty") and not getattr(self, f.name):
                continue            # extensions of the reference's schema stay off the wire unless used
            out[f.metadata.get("json", f.name)] = _plain(getattr(self, f.name))
        return out

    def to_json(self) -> str:
        return json.dumps(self.to_dict())

    @classmethod
    def from_dict(cls, d: Optional[Dict[str, Any]]):
        d = d or {}
        kwargs = {}
        for f in dataclasses.fields(cls):
            if f.metadata.get("skip"):
                continue
            key = f.metadata.get("json", f.name)
            if key not in d or d[key] is None:
                continue
            sub = f.metadata.get("cls")
            kwargs[f.name] = sub.from_dict(d[key]) if sub is not None else d[key]
        return cls(**kwargs)

    @classmethod
    def from_json(cls, s: str):
        return cls.from_dict(json.loads(s))


def _j(name: str, **kw):
    return {"json": name, **kw}


@dataclass
class TrainOptions(_Wire):
    """types.go:25-34."""
    default_parallelism: int = field(default=0, metadata=_j("default_parallelism"))
    static_parallelism: bool = field(default=False, metadata=_j("static_parallelism"))
    validate_every: int = field(default=0, metadata=_j("validate_every"))
    k: int = field(default=-1, metadata=_j("k"))
    goal_accuracy: float = field(default=100.0, metadata=_j("goal_accuracy"))
    # Extension (the reference keeps no model past the end of a job): start from the checkpoint
    # of an earlier task instead of the function's fresh initialisation.
    resume_from: str = field(default="", metadata=dict(_j("resume_from"), omitempty=True))


@dataclass
class TrainRequest(_Wire):
    """types.go:13-21."""
    model_type: str = field(default="", metadata=_j("model_type"))
    batch_size: int = field(default=64, metadata=_j("batch_size"))
    epochs: int = field(default=1, metadata=_j("epochs"))
    dataset: str = field(default="", metadata=_j("dataset"))
    lr: float = field(default=0.01, metadata=_j("lr"))
    function_name: str = field(default="", metadata=_j("function_name"))
    options: TrainOptions = field(default_factory=TrainOptions,
                                  metadata=_j("options", cls=TrainOptions))


@dataclass
class InferRequest(_Wire):
    """types.go:37-40."""
    model_id: str = field(default="", metadata=_j("model_id"))
    data: List[Any] = field(default_factory=list, metadata=_j("data"))


@dataclass
class JobState(_Wire):
    """types.go:68-71."""
    parallelism: int = field(default=0, metadata=_j("parallelism"))
    elapsed_time: float = field(default=0.0, metadata=_j("elapsed_time"))


@dataclass
class JobInfo(_Wire):
    """types.go:59-65.  Pod/Svc/Channel are process handles there; here the handle is the
    in-process job object (never serialised)."""
    id: str = field(default="", metadata=_j("id"))
    state: JobState = field(default_factory=JobState, metadata=_j("state", cls=JobState))
    handle: Any = field(default=None, metadata={"skip": True}, repr=False, compare=False)


@dataclass
class TrainTask(_Wire):
    """types.go:46-49."""
    request: TrainRequest = field(default_factory=TrainRequest,
                                  metadata=_j("request", cls=TrainRequest))
    job: JobInfo = field(default_factory=JobInfo, metadata=_j("job", cls=JobInfo))


@dataclass
class JobHistory(_Wire):
    """types.go:75-81.  epoch_duration is cumulative seconds since job start (job.go:327)."""
    validation_loss: List[float] = field(default_factory=list, metadata=_j("validation_loss"))
    accuracy: List[float] = field(default_factory=list, metadata=_j("accuracy"))
    train_loss: List[float] = field(default_factory=list, metadata=_j("train_loss"))
    parallelism: List[float] = field(default_factory=list, metadata=_j("parallelism"))
    epoch_duration: List[float] = field(default_factory=list, metadata=_j("epoch_duration"))
    # Extensions (absent in the reference; omitted from JSON when empty is not required
    # because consumers ignore unknown keys): device-timed merge latency and throughput.
    merge_seconds: List[float] = field(default_factory=list, metadata=_j("merge_seconds"))
    images_per_second: List[float] = field(default_factory=list, metadata=_j("images_per_second"))


@dataclass
class MetricUpdate(_Wire):
    """types.go:85-91 (JSON key `validations_loss` is the reference's spelling)."""
    validation_loss: float = field(default=0.0, metadata=_j("validations_loss"))
    accuracy: float = field(default=0.0, metadata=_j("accuracy"))
    train_loss: float = field(default=0.0, metadata=_j("train_loss"))
    parallelism: float = field(default=0.0, metadata=_j("parallelism"))
    epoch_duration: float = field(default=0.0, metadata=_j("epoch_duration"))
    merge_seconds: float = field(default=0.0, metadata=_j("merge_seconds"))
    images_per_second: float = field(default=0.0, metadata=_j("images_per_second"))


@dataclass
class Datapoint(_Wire):
    """types.go:94-96."""
    features: List[float] = field(default_factory=list, metadata=_j("features"))


@dataclass
class History(_Wire):
    """types.go:100-104."""
    id: str = field(default="", metadata=_j("id"))
    task: TrainRequest = field(default_factory=TrainRequest, metadata=_j("task", cls=TrainRequest))
    data: JobHistory = field(default_factory=JobHistory, metadata=_j("data", cls=JobHistory))


@dataclass
class DatasetSummary(_Wire):
    """types.go:107-111."""
    name: str = field(default="", metadata=_j("name"))
    train_set_size: int = field(default=0, metadata=_j("train_set_size"))
    test_set_size: int = field(default=0, metadata=_j("test_set_size"))


def latest_metrics(h: JobHistory) -> MetricUpdate:
    """Last element of each history series, 0 when empty (train/util.go getLatestMetrics)."""
    def last(xs):
        return float(xs[-1]) if xs else 0.0
    return MetricUpdate(validation_loss=last(h.validation_loss), accuracy=last(h.accuracy),
                        train_loss=last(h.train_loss), parallelism=last(h.parallelism),
                        epoch_duration=last(h.epoch_duration), merge_seconds=last(h.merge_seconds),
                        images_per_second=last(h.images_per_second))
