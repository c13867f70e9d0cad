"""Stand-in for redisai-py's `Client.tensorset / tensorget / close`, talking to the in-memory tensor
store of baseline/reference_arm.py over TCP (one blob per command, like AI.TENSORSET ... BLOB)."""
import pickle
import socket
import struct

import numpy as np


def _send(sock, header, blob=b""):
    h = pickle.dumps(header, protocol=pickle.HIGHEST_PROTOCOL)
    sock.sendall(struct.pack("<II", len(h), len(blob)) + h)
    if blob:
        sock.sendall(blob)


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
        arr = np.ascontiguousarray(tensor)       # redisai-py takes dtype/shape from the ndarray
        s = self._conn()
        _send(s, ("set", key, arr.dtype.str, arr.shape), arr.tobytes())
        _recv(s)
        return "OK"

    def tensorget(self, key, **kw):
        s = self._conn()
        _send(s, ("get", key))
        (status, dt, shape), blob = _recv(s)
        if status != "ok":
            raise KeyError(key)
        return np.frombuffer(blob, dtype=np.dtype(dt)).reshape(shape).copy()

    def close(self):
        if self.sock is not None:
            try:
                self.sock.close()
            finally:
                self.sock = None
