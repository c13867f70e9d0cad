"""Stand-in for redisai-py's `Client.tensorset / tensorget / close`, talking to the in-memory tensor
store of baseline/reference_arm.py over TCP (one blob per command, like AI.TENSORSET ... BLOB)."""
import pickle
import socket
import struct
import time

import numpy as np


STATS = {"get_s": 0.0, "get_n": 0, "get_bytes": 0, "set_s": 0.0, "set_n": 0, "set_bytes": 0}


def _send(sock, header, blob=b""):
    """One message = one gather-write: the 8-byte length prefix, the pickled header and the blob leave
    in a single sendmsg (looped for partial writes), so no small segment ever waits on the peer's
    delayed ACK.  Both ends also set TCP_NODELAY."""
    h = pickle.dumps(header, protocol=pickle.HIGHEST_PROTOCOL)
    parts = [memoryview(struct.pack("<II", len(h), len(blob)) + h)]
    if len(blob):
        parts.append(memoryview(blob).cast("B"))
    while parts:
        sent = sock.sendmsg(parts)
        while parts and sent >= len(parts[0]):
            sent -= len(parts[0])
            parts.pop(0)
        if parts and sent:
            parts[0] = parts[0][sent:]


def _recv_exact(sock, n):
    buf = bytearray(n)
    view = memoryview(buf)
    got = 0
    while got < n:
        r = sock.recv_into(view[got:], n - got)
        if r == 0:
            raise ConnectionError("tensor store closed the connection")
        got += r
    return buf


def _recv(sock):
    hl, bl = struct.unpack("<II", bytes(_recv_exact(sock, 8)))
    header = pickle.loads(bytes(_recv_exact(sock, hl)))
    blob = _recv_exact(sock, bl) if bl else b""
    return header, blob


class Client:
    def __init__(self, host="localhost", port=6379, **kw):
        self.addr = (host, int(port))
        self.sock = None

    def _conn(self):
        if self.sock is None:
            self.sock = socket.create_connection(self.addr)
            self.sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        return self.sock

    def tensorset(self, key, tensor, shape=None, dtype=None):
        t0 = time.perf_counter()
        arr = np.asarray(tensor)                 # redisai-py takes dtype/shape from the ndarray
        if not arr.flags.c_contiguous:
            arr = np.ascontiguousarray(arr)
        s = self._conn()
        _send(s, ("set", key, arr.dtype.str, arr.shape), arr.reshape(-1).view(np.uint8).data)
        _recv(s)
        STATS["set_s"] += time.perf_counter() - t0
        STATS["set_n"] += 1
        STATS["set_bytes"] += arr.nbytes
        return "OK"

    def tensorget(self, key, **kw):
        t0 = time.perf_counter()
        s = self._conn()
        _send(s, ("get", key))
        (status, dt, shape), blob = _recv(s)
        if status != "ok":
            raise KeyError(key)
        out = np.frombuffer(blob, dtype=np.dtype(dt)).reshape(shape)      # the receive buffer is private: no copy
        STATS["get_s"] += time.perf_counter() - t0
        STATS["get_n"] += 1
        STATS["get_bytes"] += out.nbytes
        return out

    def close(self):
        if self.sock is not None:
            try:
                self.sock.close()
            finally:
                self.sock = None
