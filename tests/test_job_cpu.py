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
