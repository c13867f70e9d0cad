"""Reference-model bookkeeping and the K-AVG average (reference: ml/pkg/model/)."""
from .store import TensorStore, weight_key, default_tensor_store
from .model import Layer, Model, ParallelSGD

__all__ = ["TensorStore", "weight_key", "default_tensor_store", "Layer", "Model", "ParallelSGD"]
