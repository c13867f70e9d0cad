"""Multi-GPU (>= 2 B200): process-per-GPU workers, peer-mapped arenas, the fused merge kernel with
its cross-rank flag barrier, against the host-side oracle."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(n, script, *args, port=29541, timeout=600, env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), script, *args]
    e = dict(os.environ)
    e["KUBEML_MERGE_TIMEOUT_S"] = "5"
    e.update(env or {})
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=e, cwd=ROOT)


def _result(proc):
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
    for line in proc.stdout.splitlines():
        if line.startswith("RESULT "):
            return json.loads(line[7:])
    raise AssertionError("no RESULT line:\n" + proc.stdout[-2000:] + proc.stderr[-2000:])


def test_merge_kernel_two_ranks_matches_average(tmp_path):
    n = min(torch.cuda.device_count(), 2)
    proc = _torchrun(n, os.path.join(ROOT, "bench", "merge_bench.py"), "--check", "--iters", "3",
                     "--sizes", "lenet,1M", port=29543)
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
    rows = [json.loads(l) for l in proc.stdout.splitlines() if l.startswith("{")]
    assert len(rows) == 2
    for r in rows:
        assert r["p2p_max_abs_err"] < 1e-6
        if "nvls_max_abs_err" in r:
            assert r["nvls_max_abs_err"] < 1e-5


def test_spmd_lenet_matches_host_oracle(tmp_path, kube_home):
    """The same job through (a) two GPU processes with the device merge and (b) the in-process host
    path (TensorStore + CPU ParallelSGD): train losses must agree."""
    n = 2
    res = _result(_torchrun(n, os.path.join(ROOT, "tests", "mp_scripts", "spmd_lenet.py"), str(tmp_path / "spmd"),
                            "2", port=29545))
    assert res["error"] is None
    dev_hist = res["history"]
    from kubeml_b200.api import TrainOptions, TrainRequest
    from kubeml_b200.controller import Cluster, make_kubeml_client
    c = Cluster(home=str(tmp_path / "host"), max_parallelism=2, options={"force_host_sync": True})
    rng = np.random.default_rng(0)
    c.shard_store.create("mnist", rng.integers(0, 255, (1024 + 64, 28, 28), dtype=np.uint8), rng.integers(0, 10, (1024 + 64,)),
                         rng.integers(0, 255, (256, 28, 28), dtype=np.uint8), rng.integers(0, 10, (256,)))
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=2, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=2, static_parallelism=True, validate_every=1, k=2))
    jid = make_kubeml_client(c).v1().networks().train(req)
    assert c.wait(jid, 300) is None
    host_hist = c.histories.get(jid).data
    c.shutdown()
    np.testing.assert_allclose(dev_hist["train_loss"], host_hist.train_loss, rtol=5e-3)
    np.testing.assert_allclose(dev_hist["validation_loss"], host_hist.validation_loss, rtol=5e-3)
    assert dev_hist["parallelism"] == [2.0, 2.0]


def test_worker_threads_on_two_gpus_of_one_process_use_the_device_merge(tmp_path):
    """The in-process Cluster runs one worker *thread* per GPU (ThreadInvoker + ThreadedDeviceSyncGroup): kernel
    attributes are per device, two GPUs launch the same kernels from one process, and the merge kernels of the two
    GPUs wait for each other.  Same job with the host merge as the oracle."""
    from kubeml_b200.api import TrainOptions, TrainRequest
    from kubeml_b200.controller import Cluster, make_kubeml_client
    hist = {}
    for mode in ("device", "host"):
        c = Cluster(home=str(tmp_path / mode), max_parallelism=2, options={"force_host_sync": mode == "host"})
        rng = np.random.default_rng(0)
        c.shard_store.create("mnist", rng.integers(0, 255, (1024, 28, 28), dtype=np.uint8), rng.integers(0, 10, (1024,)),
                             rng.integers(0, 255, (256, 28, 28), dtype=np.uint8), rng.integers(0, 10, (256,)))
        req = TrainRequest(model_type="lenet", batch_size=64, epochs=2, dataset="mnist", lr=0.01, function_name="lenet",
                           options=TrainOptions(default_parallelism=2, static_parallelism=True, validate_every=1, k=2))
        jid = make_kubeml_client(c).v1().networks().train(req)
        assert c.wait(jid, 300) is None
        assert c.jobs[jid].device_mode == (mode == "device")
        hist[mode] = c.histories.get(jid).data
        c.shutdown()
    assert hist["device"].parallelism == [2.0, 2.0]
    np.testing.assert_allclose(hist["device"].train_loss, hist["host"].train_loss, rtol=5e-3)
    np.testing.assert_allclose(hist["device"].validation_loss, hist["host"].validation_loss, rtol=5e-3)


def test_elastic_parallelism_grows_between_epochs(tmp_path):
    """Scheduler policy: 1 worker, then +1 on second sight; the joining rank maps the peers' arenas,
    copies the merged model and takes part in the next epoch's merges."""
    res = _result(_torchrun(2, os.path.join(ROOT, "tests", "mp_scripts", "spmd_elastic.py"), str(tmp_path), port=29547))
    assert res["error"] is None
    h = res["history"]
    assert h["parallelism"][0] == 1.0 and h["parallelism"][1] == 2.0
    assert len(h["train_loss"]) == 3 and all(np.isfinite(h["train_loss"]))
    assert len(h["accuracy"]) == 3


def _ngpus():
    return torch.cuda.device_count()


@pytest.mark.parametrize("n", [2, 4, 8])
def test_failed_worker_does_not_stall_or_count_in_the_device_merge(tmp_path, n):
    """Function 1 raises at the start of round 1 of epoch 1.  The reference lets the epoch stand when one
    function answered and divides by the replicas it received (ml/pkg/train/job.go:383-398,
    train/util.go:144-166); on the device path the failing rank marks itself dead in every signal pad, so the
    survivors neither wait for it (no timeout is involved) nor count it, and it is back in epoch 2."""
    if _ngpus() < n:
        pytest.skip(f"needs {n} GPUs")
    import time
    t0 = time.time()
    res = _result(_torchrun(n, os.path.join(ROOT, "tests", "mp_scripts", "spmd_fault.py"), str(tmp_path), "1@1:1",
                            port=29551, env={"KUBEML_MERGE_TIMEOUT_S": "30"}))
    assert res["error"] is None
    h = res["history"]
    assert len(h["train_loss"]) == 2 and all(np.isfinite(h["train_loss"]))
    assert h["parallelism"] == [float(n), float(n)]
    assert res["recovered"] == 0                      # nobody had to time out
    assert time.time() - t0 < 120                     # far below n_rounds x 30 s of barrier timeouts


def test_stalled_worker_is_dropped_after_the_round_timeout_and_evicted(tmp_path):
    """Function 1 goes quiet for 4 s in round 1 (no failure announced).  The survivors' merge times out after
    0.5 s without touching anything, they record the rank as dead, merge again without it and finish the
    epoch; the straggler finds itself evicted when it arrives and fails instead of corrupting the model."""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    res = _result(_torchrun(2, os.path.join(ROOT, "tests", "mp_scripts", "spmd_fault.py"), str(tmp_path),
                            "1@1:1/stall=4", port=29553, env={"KUBEML_MERGE_TIMEOUT_MS": "500"}))
    assert res["error"] is None
    h = res["history"]
    assert len(h["train_loss"]) == 2 and all(np.isfinite(h["train_loss"]))
    assert res["recovered"] >= 1


def test_two_jobs_train_side_by_side_on_disjoint_gpu_pairs(tmp_path):
    """The reference runs one pod per job, several at once (ml/pkg/ps/api.go:139-222).  Here two jobs take two GPUs
    each out of the same worker pool and train concurrently: separate arenas, peer mappings, multicast objects
    and signal pads per job, fan-outs through per-rank mailboxes."""
    if _ngpus() < 4:
        pytest.skip("needs 4 GPUs")
    res = _result(_torchrun(4, os.path.join(ROOT, "tests", "mp_scripts", "spmd_two_jobs.py"), str(tmp_path), port=29555))
    a, b = res["jobA"], res["jobB"]
    assert a["error"] is None and b["error"] is None
    assert sorted(a["ranks"] + b["ranks"]) == [0, 1, 2, 3] and not set(a["ranks"]) & set(b["ranks"])
    assert len(a["train_loss"]) == 3 and len(b["train_loss"]) == 3
    assert all(np.isfinite(a["train_loss"])) and all(np.isfinite(b["train_loss"]))
    assert a["train_loss"] != b["train_loss"]            # different learning rates: really two models
    assert res["overlap_s"] > 0.2                         # they ran at the same time, not one after the other
