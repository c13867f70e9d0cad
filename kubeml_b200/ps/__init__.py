"""Parameter-server role: job registry, job factory, metrics (reference: ml/pkg/ps/)."""
from .metrics import MetricsRegistry
from .parameter_server import ParameterServer

__all__ = ["MetricsRegistry", "ParameterServer"]
