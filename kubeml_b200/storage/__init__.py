"""Dataset storage: the replacement for MongoDB + the Flask storage service
(python/storage/api.py)."""
from .store import ShardStore, default_store, set_default_store
from .service import StorageService

__all__ = ["ShardStore", "default_store", "set_default_store", "StorageService"]
