"""Stand-in for the three MongoDB calls of the reference dataset
(python/kubeml/kubeml/dataset.py:97-123,196-203): list_database_names, count_documents, find by
`_id` range.  Documents are the ones the reference storage service writes: pickled 64-sample
chunks (python/storage/utils.py:17-25)."""
from . import errors  # noqa: F401

DATABASES = {}     # name -> {"train": [doc, ...], "test": [doc, ...]}; filled by the benchmark arm


class _Collection:
    def __init__(self, docs):
        self.docs = docs

    def count_documents(self, flt):
        return len(self.docs)

    def find(self, flt):
        rng = flt.get("_id", {})
        lo, hi = rng.get("$gte", 0), rng.get("$lte", len(self.docs) - 1)
        return iter(self.docs[lo:hi + 1])


class _Database:
    def __init__(self, cols):
        self.cols = cols

    def __getitem__(self, name):
        return _Collection(self.cols.get(name, []))


class MongoClient:
    def __init__(self, host=None, port=None, **kw):
        pass

    def list_database_names(self):
        return list(DATABASES.keys())

    def __getitem__(self, name):
        return _Database(DATABASES.get(name, {}))

    def close(self):
        pass
