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


def test_dead_worker_is_detected_by_heartbeat_not_by_a_collective_timeout(tmp_path):
    """worker liveness.  Rank 1 dies without a word while rank 0 waits for its result; the fan-out
    must come back with an error for that function after the heartbeat timeout (2 s here), not after the
    10-minute process-group timeout, and the next fan-out must not wait for the dead rank at all."""
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    script = os.path.join(ROOT, "tests", "mp_scripts", "spmd_liveness.py")
    procs = []
    for r in range(2):
        env = dict(os.environ, CUDA_VISIBLE_DEVICES="", RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), KUBEML_HEARTBEAT_TIMEOUT_S="2",
                   KUBEML_HOME=str(tmp_path), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, script], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                                      text=True, env=env, cwd=ROOT))
    out0, err0 = procs[0].communicate(timeout=180)
    procs[1].wait(timeout=60)
    assert procs[1].returncode == 17
    assert procs[0].returncode == 0, err0[-3000:]
    res = next(json.loads(l[7:]) for l in out0.splitlines() if l.startswith("RESULT "))
    assert [p[0]["rank"] for p in res["first"]] == [0, 1]
    dead = res["after_death"][1]
    assert dead[1] == 503 and "not responding" in dead[0]["error"]
    assert res["after_death"][0][1] == 200                 # rank 0's own share still ran
    assert 1.0 < res["detect_s"] < 20.0
    assert res["later"][1][1] == 503 and res["later_s"] < 1.0


def test_fan_outs_to_disjoint_ranks_overlap(tmp_path):
    """Per-rank mailboxes: two fan-outs (two jobs on disjoint GPU subsets) do not serialise behind one lock, and
    the rank allocator hands out disjoint subsets."""
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    script = os.path.join(ROOT, "tests", "mp_scripts", "spmd_mailbox.py")
    procs = []
    for r in range(3):
        env = dict(os.environ, CUDA_VISIBLE_DEVICES="", RANK=str(r), WORLD_SIZE="3", LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), KUBEML_HOME=str(tmp_path), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, script], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                                      text=True, env=env, cwd=ROOT))
    out0, err0 = procs[0].communicate(timeout=180)
    for p in procs[1:]:
        p.wait(timeout=60)
    assert procs[0].returncode == 0, err0[-3000:]
    res = next(json.loads(l[7:]) for l in out0.splitlines() if l.startswith("RESULT "))
    assert res["both_s"] < 1.8                                    # 2 x 1 s of work in ~1 s
    assert res["res"]["1"][0][0]["rank"] == 1 and res["res"]["2"][0][0]["rank"] == 2
    assert res["res"]["1"][0][0]["members"] == [1]
    assert res["held"] == [0, 1] and res["refused_when_short"] is True
    assert [p[0]["rank"] for p in res["everyone"]] == [0, 1, 2]
