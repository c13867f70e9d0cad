"""Addresses and defaults of the single-host control plane.

The reference hard-codes Kubernetes service DNS names (ml/pkg/api/const.go:4-14) and debug
NodePort addresses (:19-30).  On one 8xB200 host every role lives on 127.0.0.1, so the only
constants left are ports and the parallelism defaults (const.go:16, :25).
"""
import os

HOST = os.environ.get("KUBEML_HOST", "127.0.0.1")
CONTROLLER_PORT = int(os.environ.get("KUBEML_CONTROLLER_PORT", "10100"))
SCHEDULER_PORT = int(os.environ.get("KUBEML_SCHEDULER_PORT", "10200"))
PARAMETER_SERVER_PORT = int(os.environ.get("KUBEML_PS_PORT", "10300"))
STORAGE_PORT = int(os.environ.get("KUBEML_STORAGE_PORT", "10400"))
METRICS_PORT = int(os.environ.get("KUBEML_METRICS_PORT", "8080"))

CONTROLLER_URL = f"http://{HOST}:{CONTROLLER_PORT}"

# ml/pkg/api/const.go:16 / :25
DEFAULT_PARALLELISM = 5
DEBUG_PARALLELISM = 2

# Number of samples per stored subset ("document"); python/kubeml/kubeml/util.py:10 and
# python/storage/api.py:135 agree on 64.
STORAGE_SUBSET_SIZE = 64

# A single B200 host exposes at most this many workers; the reference never clamps
# (scheduler/policy.go:68-89), we do.
MAX_GPUS = 8


def home() -> str:
    """Root directory for datasets, histories, checkpoints and function sources."""
    return os.environ.get("KUBEML_HOME", os.path.join(os.path.expanduser("~"), ".kubeml_b200"))
