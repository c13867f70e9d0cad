"""kubeml_b200 — a Blackwell-native training runtime with KubeML's capabilities.

User-facing surface (same names as the reference `kubeml` package,
python/kubeml/kubeml/__init__.py): `KubeModel`, `KubeDataset`.
"""
from .dataset import KubeDataset
from .network import KubeModel
from .exceptions import (KubeMLException, MergeError, DataError, InvalidFormatError, StorageError,
                         DatasetNotFoundError, InvalidArgsError)

__version__ = "0.1.0"
__all__ = ["KubeModel", "KubeDataset", "KubeMLException", "MergeError", "DataError",
           "InvalidFormatError", "StorageError", "DatasetNotFoundError", "InvalidArgsError"]
