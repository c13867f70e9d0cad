"""Storage service role: `POST /dataset/{name}` (4 multipart files), `DELETE /dataset/{name}`,
`GET /health` — the routes of python/storage/api.py:37-51, backed by the local ShardStore."""
from __future__ import annotations

import os
import tempfile
import uuid
from typing import Dict, Optional, Tuple

from ..exceptions import KubeMLException
from ..httpd import Request, Router, Server, decode_multipart
from .store import ShardStore, default_store


class StorageService:
    def __init__(self, store: Optional[ShardStore] = None):
        self.store = store or default_store()

    # ---- operations (in-process API) ---------------------------------------------------
    def upload(self, name: str, files: Dict[str, Tuple[str, bytes]]) -> dict:
        """`files`: field -> (filename, bytes).  Mirrors upload_dataset/_process_datasets
        (api.py:58-142): reject empty requests and duplicates with 400, spill each upload under
        an 8-char upload id so concurrent uploads cannot clobber each other, then ingest."""
        if not files:
            raise KubeMLException("Request does not include a file", 400)
        if self.store.exists(name):
            raise KubeMLException(f"Dataset {name} already exists", 400)
        upload_id = uuid.uuid4().hex[:8]
        tmpdir = tempfile.mkdtemp(prefix=f"kubeml-upload-{upload_id}-")
        paths = {}
        try:
            for field in ("x-train", "y-train", "x-test", "y-test"):
                if field not in files:
                    raise KubeMLException("Request does not include a file", 400)
                filename, content = files[field]
                ext = filename.rsplit(".", 1)[-1].lower() if "." in filename else ""
                if ext not in ("npy", "pkl"):
                    raise KubeMLException(
                        "File extension not supported, must be one of [npy, pkl]", 400)
                p = os.path.join(tmpdir, f"{field}-{upload_id}.{ext}")
                with open(p, "wb") as f:
                    f.write(content)
                paths[field] = p
            self.store.create_from_files(name, paths)
        finally:
            for p in paths.values():
                try:
                    os.remove(p)
                except OSError:
                    pass
            try:
                os.rmdir(tmpdir)
            except OSError:
                pass
        return {"result": "Dataset created"}

    def delete(self, name: str) -> dict:
        self.store.delete(name)  # 404 "Dataset does not exist" (api.py:145-156)
        return {"result": "Dataset deleted"}

    # ---- HTTP --------------------------------------------------------------------------
    def router(self) -> Router:
        r = Router()
        r.add("GET", "/health", lambda req: (200, None))
        r.add("POST", "/dataset/{name}", self._http_upload)
        r.add("DELETE", "/dataset/{name}", lambda req: self.delete(req.params["name"]))
        return r

    def _http_upload(self, req: Request):
        ctype = req.headers.get("Content-Type") or req.headers.get("content-type") or ""
        return self.upload(req.params["name"], decode_multipart(req.body, ctype))

    def serve(self, host: str, port: int) -> Server:
        return Server(self.router(), host, port).start()
