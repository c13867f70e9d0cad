"""Scheduler role: FIFO queue + parallelism policy (reference: ml/pkg/scheduler/)."""
from .policy import ThroughputBasedPolicy, TaskOperation, SchedulerPolicy
from .queue import SchedulerQueue
from .scheduler import Scheduler

__all__ = ["ThroughputBasedPolicy", "TaskOperation", "SchedulerPolicy", "SchedulerQueue", "Scheduler"]
