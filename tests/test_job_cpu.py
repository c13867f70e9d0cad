"""Train-job lifecycle on the CPU (LeNet, synthetic MNIST-shape data,
parallelism 2, K=16, batch 64) against a single-process restatement of K-AVG; plus elastic
parallelism, fault tolerance, stop and goal accuracy."""
import copy
import os

import numpy as np
import pytest
import torch

from kubeml_b200.api import TrainOptions, TrainRequest
from kubeml_b200.util import get_subset_period, split_minibatches


def _mnist(store, n=64 * 40, nt=64 * 4, seed=0):
    rng = np.random.default_rng(seed)
    store.create("mnist", rng.integers(0, 255, (n, 28, 28), dtype=np.uint8), rng.integers(0, 10, (n,)),
                 rng.integers(0, 255, (nt, 28, 28), dtype=np.uint8), rng.integers(0, 10, (nt,)))


def _run(cluster, req, timeout=600):
    from kubeml_b200.controller import make_kubeml_client
    cl = make_kubeml_client(cluster)
    jid = cl.v1().networks().train(req)
    err = cluster.wait(jid, timeout)
    return jid, err, cluster.jobs[jid]


@pytest.fixture()
def cluster(kube_home, monkeypatch):
    from kubeml_b200.controller import Cluster
    monkeypatch.setenv("KUBEML_CUDA_GRAPH", "0")
    c = Cluster(home=kube_home, max_parallelism=4, options={"force_host_sync": True})
    yield c
    c.shutdown()


def _oracle_kavg(store, N, K, B, lr, epochs):
    """K-AVG exactly as SURVEY 3.7 states it, in one process with plain torch."""
    from kubeml_b200.functions.function_lenet import LeNet
    torch.manual_seed(42)
    ref = LeNet()
    x_all, y_all = store.arrays("mnist", "train")
    D = store.count_docs("mnist", "train")
    losses = []
    for _ in range(epochs):
        shards = split_minibatches(range(D), N)
        per = [get_subset_period(K, B, s) for s in shards]
        starts = [list(range(s.start, s.stop, p)) for s, p in zip(shards, per)]
        nrounds = max(len(s) for s in starts)
        loss_sum, nb = [0.0] * N, [0] * N
        for rnd in range(nrounds):
            replicas = []
            for r in range(N):
                if rnd >= len(starts[r]):
                    continue
                net = copy.deepcopy(ref)
                opt = torch.optim.SGD(net.parameters(), lr=lr, momentum=0.9, weight_decay=1e-4)
                lo, hi = starts[r][rnd] * 64, min(shards[r].stop, starts[r][rnd] + per[r]) * 64
                x = torch.from_numpy(np.array(x_all[lo:hi])).float().div(255).sub(0.1307).div(0.3081).view(-1, 1, 28, 28)
                y = torch.from_numpy(np.array(y_all[lo:hi])).long()
                for i in range(0, len(x), B):
                    opt.zero_grad()
                    loss = torch.nn.functional.cross_entropy(net(x[i:i + B]), y[i:i + B])
                    loss.backward(); opt.step()
                    loss_sum[r] += loss.item(); nb[r] += 1
                replicas.append(net.state_dict())
            with torch.no_grad():
                for k in ref.state_dict():
                    acc = replicas[0][k].clone()
                    for sd in replicas[1:]:
                        acc += sd[k]
                    ref.state_dict()[k].copy_(acc / len(replicas))
        losses.append(float(np.mean([loss_sum[r] / nb[r] for r in range(N)])))
    return ref.state_dict(), losses


def test_config1_lenet_matches_kavg_oracle(cluster):
    torch.set_num_threads(1)
    _mnist(cluster.shard_store)
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=2, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=2, static_parallelism=True, validate_every=1, k=16))
    jid, err, job = _run(cluster, req)
    assert err is None
    h = job.history
    assert len(h.train_loss) == 2 and len(h.accuracy) == 2 and h.parallelism == [2.0, 2.0]
    assert h.epoch_duration[1] >= h.epoch_duration[0]                  # cumulative since job start
    # 40 docs / 2 workers = 20 docs each, 16 docs per round -> 2 rounds per epoch
    assert [len(m["participants"]) for m in job.merge_log] == [2, 2, 2, 2]
    want_sd, want_losses = _oracle_kavg(cluster.shard_store, 2, 16, 64, 0.01, 2)
    np.testing.assert_allclose(h.train_loss, want_losses, rtol=1e-4)
    for k, v in want_sd.items():
        got = cluster.tensor_store.tensorget(f"{jid}:{k}")             # the reference model is kept
        np.testing.assert_allclose(got, v.numpy(), rtol=1e-4, atol=1e-6)
    assert cluster.tensor_store.keys(f"{jid}:*/*") == []               # replicas are dropped
    # history persisted and a checkpoint written
    assert cluster.histories.get(jid).data.train_loss == h.train_loss
    from kubeml_b200.utils.checkpoint import load_checkpoint
    sd, hist = load_checkpoint(cluster.checkpoint_dir, jid)
    assert list(sd.keys()) == list(want_sd.keys()) and hist.id == jid


def test_uneven_shards_shrink_the_participant_set(cluster):
    _mnist(cluster.shard_store, n=64 * 7)                              # 7 docs over 3 workers: 3,2,2
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=1, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=3, static_parallelism=True, k=1))
    _, err, job = _run(cluster, req)
    assert err is None
    assert [sorted(m["participants"]) for m in job.merge_log] == [[0, 1, 2], [0, 1, 2], [0]]


def test_dynamic_parallelism_follows_the_policy(cluster):
    _mnist(cluster.shard_store, n=64 * 12)
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=3, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=2, static_parallelism=False, k=-1))
    _, err, job = _run(cluster, req)
    assert err is None
    p = job.history.parallelism
    assert p[0] == 2.0 and p[1] == 3.0                                 # second sight: always +1
    assert 1.0 <= p[2] <= 4.0 and abs(p[2] - p[1]) <= 1.0              # then +-1 or keep, clamped to 4


def test_limit_parallelism_env_freezes_n(cluster, monkeypatch):
    monkeypatch.setenv("LIMIT_PARALLELISM", "true")
    _mnist(cluster.shard_store, n=64 * 8)
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=2, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=2, static_parallelism=False, k=-1))
    _, err, job = _run(cluster, req)
    assert err is None and job.history.parallelism == [2.0, 2.0]


def test_injected_fault_is_tolerated_and_excluded(kube_home, monkeypatch):
    from kubeml_b200.controller import Cluster
    monkeypatch.setenv("KUBEML_CUDA_GRAPH", "0")
    fault = lambda func_id, epoch, rnd, idx: func_id == 1 and epoch == 1 and rnd == 0 and idx == 0
    c = Cluster(home=kube_home, max_parallelism=4, options={"force_host_sync": True, "inject_fault": fault})
    _mnist(c.shard_store, n=64 * 8)
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=2, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=2, static_parallelism=True, k=-1))
    _, err, job = _run(c, req)
    assert err is None                                                 # >= 1 function answered: epoch stands
    assert [sorted(m["participants"]) for m in job.merge_log] == [[0], [0, 1]]
    assert len(job.history.train_loss) == 2
    c.shutdown()


def test_all_functions_failing_fails_the_job(kube_home, monkeypatch):
    from kubeml_b200.controller import Cluster
    c = Cluster(home=kube_home, max_parallelism=2, options={"force_host_sync": True, "inject_fault": lambda *a: True})
    _mnist(c.shard_store, n=64 * 4)
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=1, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=2, static_parallelism=True, k=-1))
    _, err, job = _run(c, req)
    assert err is not None and "injected fault" in err
    assert c.ps.list_tasks() == []                                     # finished jobs leave the index
    c.shutdown()


def test_goal_accuracy_stops_early_and_stop_request(cluster):
    _mnist(cluster.shard_store, n=64 * 4)
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=5, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=1, static_parallelism=True, validate_every=1, k=-1,
                                            goal_accuracy=0.0))
    _, err, job = _run(cluster, req)
    assert err is None and len(job.history.train_loss) == 1 and len(job.history.accuracy) == 1


def test_missing_dataset_and_function_errors(cluster):
    req = TrainRequest(model_type="x", batch_size=64, epochs=1, dataset="nope", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=1, static_parallelism=True))
    _, err, _ = _run(cluster, req)
    assert err is not None and "Dataset not found" in err
    req.function_name = "does_not_exist"
    _, err, _ = _run(cluster, req)
    assert err is not None and "does not exist" in err


def test_resume_from_checkpoint_continues_where_the_first_task_stopped(cluster):
    """W24: a task started with `resume_from` trains on from the checkpointed model: one epoch + one resumed
    epoch give the same model as two epochs in one task (same data order, optimizer state is reset at every
    sync anyway)."""
    torch.set_num_threads(1)
    _mnist(cluster.shard_store)
    mk = lambda epochs, resume="": TrainRequest(
        model_type="lenet", batch_size=64, epochs=epochs, dataset="mnist", lr=0.01, function_name="lenet",
        options=TrainOptions(default_parallelism=2, static_parallelism=True, validate_every=0, k=16, resume_from=resume))
    j1, err, _ = _run(cluster, mk(1))
    assert err is None
    j2, err, job2 = _run(cluster, mk(1, resume=j1))
    assert err is None
    want_sd, _ = _oracle_kavg(cluster.shard_store, 2, 16, 64, 0.01, 2)
    for k, v in want_sd.items():
        got = cluster.tensor_store.tensorget(f"{j2}:{k}")
        np.testing.assert_allclose(got, v.numpy(), rtol=1e-4, atol=1e-6)
    _, err, _ = _run(cluster, mk(1, resume="nosuchjob"))
    assert err is not None and "no checkpoint" in str(err)


def test_standalone_job_runs_in_its_own_process_and_reports_over_http(kube_home, monkeypatch):
    """STANDALONE_JOBS (ml/cmd/ml/main.go:119-129, ml/pkg/ps/api.go:162-207): the parameter server starts
    `python -m kubeml_b200 --jobPort P --jobId ID`, waits for /health, posts the task to /start; the job process
    sends its metrics and its end back to the parameter server, asks the scheduler for its parallelism, and its
    output is the job log.  Same losses as the job run as a thread."""
    from kubeml_b200.controller import Cluster, make_kubeml_client
    monkeypatch.setenv("KUBEML_CUDA_GRAPH", "0")
    monkeypatch.setenv("CUDA_VISIBLE_DEVICES", "")
    monkeypatch.setenv("KUBEML_JOB_READY_TIMEOUT_S", "180")     # a cold `import torch` in the child can take a minute
    # two epochs: the policy's first two answers (default parallelism, then +1) do not depend on epoch times
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=2, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=2, static_parallelism=False, validate_every=1, k=4))
    hist = {}
    for mode in ("thread", "process"):
        monkeypatch.setenv("STANDALONE_JOBS", "true" if mode == "process" else "false")
        c = Cluster(home=os.path.join(kube_home, mode), max_parallelism=4, options={"force_host_sync": True})
        assert c.ps.standalone_jobs == (mode == "process")
        _mnist(c.shard_store, n=64 * 16)
        seen = []
        orig = c.ps.update_metrics
        c.ps.update_metrics = lambda jid, m, _o=orig: (seen.append(m), _o(jid, m))[1]
        monkeypatch.setenv("KUBEML_HOME", os.path.join(kube_home, mode))
        jid = make_kubeml_client(c).v1().networks().train(req)
        err = c.wait(jid, 300)
        assert err is None, err
        hist[mode] = c.histories.get(jid).data
        if mode == "process":
            assert jid not in c.jobs                              # nothing of the job ran in this process
            assert len(seen) >= 2                                 # MetricUpdates of every epoch, over HTTP
            log = open(os.path.join(kube_home, mode, "logs", f"job-{jid}.log")).read()
            assert "Starting train job" in log and "serving on" in log
            assert c.ps.list_tasks() == []
        c.shutdown()
    assert hist["process"].parallelism == hist["thread"].parallelism == [2.0, 3.0]    # the policy was asked over HTTP
    np.testing.assert_allclose(hist["process"].train_loss, hist["thread"].train_loss, rtol=1e-5)
    np.testing.assert_allclose(hist["process"].accuracy, hist["thread"].accuracy, rtol=1e-5)


def test_standalone_job_that_dies_is_reported_failed(kube_home, monkeypatch):
    """Liveness of a job process: killed in mid-run, it becomes a failed task instead of a task that is listed
    for ever (the reference relies on the pod's readiness probe and phase, ml/pkg/ps/job_pod.go:120-131,172-186)."""
    import time
    from kubeml_b200.controller import Cluster, make_kubeml_client
    monkeypatch.setenv("KUBEML_CUDA_GRAPH", "0")
    monkeypatch.setenv("CUDA_VISIBLE_DEVICES", "")
    monkeypatch.setenv("KUBEML_JOB_READY_TIMEOUT_S", "180")     # a cold `import torch` in the child can take a minute
    monkeypatch.setenv("STANDALONE_JOBS", "true")
    c = Cluster(home=kube_home, max_parallelism=2, options={"force_host_sync": True})
    _mnist(c.shard_store, n=64 * 64)
    req = TrainRequest(model_type="lenet", batch_size=16, epochs=50, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=2, static_parallelism=True, validate_every=0, k=1))
    jid = make_kubeml_client(c).v1().networks().train(req)
    deadline = time.time() + 60
    while time.time() < deadline and not any(t.job.id == jid and t.job.handle is not None for t in c.ps.list_tasks()):
        time.sleep(0.05)
    task = next(t for t in c.ps.list_tasks() if t.job.id == jid)
    task.job.handle.proc.kill()
    err = c.wait(jid, 60)
    assert err is not None and "without reporting" in err
    assert c.ps.list_tasks() == []
    c.shutdown()


def test_inference_is_served_from_the_checkpoint_of_a_finished_task(cluster):
    """`kubeml infer --id <task>` after the task ended (the reference's inference only half exists):
    the task's own function answers from the checkpoint the task left; while it runs, from the live model."""
    from kubeml_b200.api import InferRequest
    from kubeml_b200.controller import make_kubeml_client
    from kubeml_b200.exceptions import KubeMLException
    _mnist(cluster.shard_store, n=64 * 8)
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=1, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=2, static_parallelism=True, validate_every=0, k=2))
    jid, err, job = _run(cluster, req)
    assert err is None and cluster.ps.list_tasks() == []
    cl = make_kubeml_client(cluster)
    x = np.random.default_rng(1).integers(0, 255, (5, 28, 28)).astype(np.float32).tolist()
    preds = cl.v1().networks().infer(InferRequest(model_id=jid, data=x))
    assert len(preds) == 5 and all(0 <= int(p) < 10 for p in preds)
    # same answer as the trained network evaluated directly
    from kubeml_b200.functions.function_lenet import LeNet
    from kubeml_b200.utils.checkpoint import load_checkpoint
    sd, _ = load_checkpoint(cluster.checkpoint_dir, jid)
    net = LeNet(); net.load_state_dict(sd); net.eval()
    with torch.no_grad():
        want = net(torch.tensor(x).view(-1, 1, 28, 28)).argmax(1).tolist()
    assert [int(p) for p in preds] == want
    assert cl.v1().networks().infer(InferRequest(model_id=jid, data=x[:2])) == preds[:2]      # the warm worker again
    with pytest.raises(KubeMLException) as e:
        cl.v1().networks().infer(InferRequest(model_id="deadbeef", data=x))
    assert e.value.status_code == 404
