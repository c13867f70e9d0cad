"""Controller role + typed client (reference: ml/pkg/controller/, ml/pkg/controller/client/)."""
from .controller import Controller
from .client import KubemlClient, make_kubeml_client
from .cluster import Cluster

__all__ = ["Controller", "KubemlClient", "make_kubeml_client", "Cluster"]
