"""File-descriptor exchange between the per-GPU worker processes.

A VMM allocation (or a multicast object) is exported as a POSIX fd; the only way to hand an fd to
another process is SCM_RIGHTS over a Unix-domain socket.  Every rank listens on
`<dir>/rank<r>.sock`, connects to each peer, and sends `(rank, tag, fds)`.  This replaces the
reference's RedisAI connection set-up (python/kubeml/kubeml/network.py:19-26,52): the "address" of a
replica is now a mapped pointer instead of a Redis key.
"""
from __future__ import annotations

import os
import socket
import struct
import time
from typing import Dict, List, Optional


def default_socket_dir() -> str:
    d = os.environ.get("KUBEML_SOCK_DIR")
    if not d:
        port = os.environ.get("MASTER_PORT", "0")
        d = f"/tmp/kubeml_b200_{os.getuid()}_{port}"
    os.makedirs(d, exist_ok=True)
    return d


class FdExchange:
    def __init__(self, rank: int, world: int, sock_dir: Optional[str] = None):
        self.rank, self.world = rank, world
        self.dir = sock_dir or default_socket_dir()
        self.path = os.path.join(self.dir, f"rank{rank}.sock")
        try:
            os.unlink(self.path)
        except FileNotFoundError:
            pass
        self.server = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        self.server.bind(self.path)
        self.server.listen(world + 4)
        self._pending: Dict[tuple, List[int]] = {}

    def _connect(self, peer: int, timeout: float) -> socket.socket:
        path = os.path.join(self.dir, f"rank{peer}.sock")
        deadline = time.monotonic() + timeout
        while True:
            s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            try:
                s.connect(path)
                return s
            except (FileNotFoundError, ConnectionRefusedError):
                s.close()
                if time.monotonic() > deadline:
                    raise TimeoutError(f"rank {self.rank}: peer {peer} never opened {path}")
                time.sleep(0.01)

    def all_gather_fds(self, fds: List[int], tag: int, timeout: float = 120.0,
                       members: Optional[List[int]] = None) -> List[List[int]]:
        """Every rank of `members` (default: the whole world) contributes `fds`; returns, per member in the
        order of `members`, the fds it contributed (received copies are new descriptors owned by the caller;
        the own entry is the list passed in).  Exchanges of disjoint member sets (two jobs on different GPU
        subsets) do not interfere: messages are matched by (sender, tag) and tags are scoped per job."""
        members = list(range(self.world)) if members is None else list(members)
        for peer in members:
            if peer == self.rank:
                continue
            s = self._connect(peer, timeout)
            try:
                # an fd in flight is kept alive by the kernel, so no acknowledgement is needed
                socket.send_fds(s, [struct.pack("<iii", self.rank, tag, len(fds))], fds)
            finally:
                s.close()
        out: Dict[int, Optional[List[int]]] = {m: None for m in members}
        out[self.rank] = list(fds)
        for key in [k for k in self._pending if k[1] == tag]:
            if key[0] in out:
                out[key[0]] = self._pending.pop(key)
        self.server.settimeout(timeout)
        while any(o is None for o in out.values()):
            conn, _ = self.server.accept()
            try:
                msg, got, _, _ = socket.recv_fds(conn, 12, 64)
                src, t, n = struct.unpack("<iii", msg[:12])
            finally:
                conn.close()
            got = list(got)[:n] if n else []
            if t == tag and src in out:
                out[src] = got
            else:
                self._pending[(src, t)] = got
        return [out[m] for m in members]  # type: ignore[misc]

    def broadcast_fds(self, fds: List[int], root: int, tag: int, timeout: float = 120.0,
                      members: Optional[List[int]] = None) -> List[int]:
        members = list(range(self.world)) if members is None else list(members)
        got = self.all_gather_fds(fds if self.rank == root else [], tag, timeout, members)
        return got[members.index(root)]

    def all_agree(self, ok: bool, tag: int, timeout: float = 120.0, members: Optional[List[int]] = None) -> bool:
        """True iff every member passed ok=True (a yes vote travels as one descriptor)."""
        members = list(range(self.world)) if members is None else list(members)
        fd = os.open(os.devnull, os.O_RDONLY) if ok else -1
        got = self.all_gather_fds([fd] if ok else [], tag, timeout, members)
        yes = all(len(g) > 0 for g in got)
        for m, g in zip(members, got):
            for f in g:
                if m != self.rank or f == fd:
                    os.close(f)
        return yes

    def barrier(self, tag: int, timeout: float = 120.0, members: Optional[List[int]] = None) -> None:
        self.all_gather_fds([], tag, timeout, members)

    def close(self) -> None:
        try:
            self.server.close()
            os.unlink(self.path)
        except OSError:
            pass
