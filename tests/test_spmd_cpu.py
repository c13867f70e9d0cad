"""Multi-process host-side logic on the CPU: torchrun + gloo, world size 2."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_spmd_collective_sync_two_processes(tmp_path):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29571", os.path.join(ROOT, "tests", "mp_scripts", "spmd_cpu_collective.py"),
           str(tmp_path)]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    proc = subprocess.run(cmd, capture_output=True, text=True, timeout=420, env=env, cwd=ROOT)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-3000:]
    res = next(json.loads(l[7:]) for l in proc.stdout.splitlines() if l.startswith("RESULT "))
    assert res["error"] is None
    h = res["history"]
    assert len(h["train_loss"]) == 2 and len(h["accuracy"]) == 2 and h["parallelism"] == [2.0, 2.0]
    assert all(np.isfinite(h["train_loss"]))


def test_fd_exchange_between_processes(tmp_path):
    code = r"""
import os, sys
sys.path.insert(0, %r)
from kubeml_b200.parallel.peer import FdExchange
rank, world, d = int(sys.argv[1]), 3, sys.argv[2]
x = FdExchange(rank, world, d)
path = os.path.join(d, f"payload{rank}")
open(path, "w").write(f"hello from {rank}")
fd = os.open(path, os.O_RDONLY)
got = x.all_gather_fds([fd], tag=1)
assert x.all_agree(True, tag=2) is True
assert x.all_agree(rank != 1, tag=3) is False
for p in range(world):
    if p != rank:
        assert os.pread(got[p][0], 100, 0) == f"hello from {p}".encode()
x.barrier(tag=4)
x.close()
print("ok", rank)
""" % ROOT
    procs = [subprocess.Popen([sys.executable, "-c", code, str(r), str(tmp_path)], stdout=subprocess.PIPE,
                              stderr=subprocess.PIPE, text=True) for r in range(3)]
    for r, p in enumerate(procs):
        out, err = p.communicate(timeout=120)
        assert p.returncode == 0 and f"ok {r}" in out, err[-2000:]
