"""Wire schema shared by every control-plane role (reference: ml/pkg/api/types.go:13-111)."""
from .types import (TrainRequest, TrainOptions, InferRequest, TrainTask, JobInfo, JobState,
                    JobHistory, MetricUpdate, History, DatasetSummary, Datapoint)
from . import const

__all__ = ["TrainRequest", "TrainOptions", "InferRequest", "TrainTask", "JobInfo", "JobState",
           "JobHistory", "MetricUpdate", "History", "DatasetSummary", "Datapoint", "const"]
