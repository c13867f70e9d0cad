"""The built-in example functions run end to end on the CPU (tiny synthetic datasets)."""
import numpy as np
import pytest

from kubeml_b200.api import TrainOptions, TrainRequest


@pytest.fixture()
def cluster(kube_home, monkeypatch):
    from kubeml_b200.controller import Cluster
    monkeypatch.setenv("KUBEML_CUDA_GRAPH", "0")
    c = Cluster(home=kube_home, max_parallelism=2, options={"force_host_sync": True})
    yield c
    c.shutdown()


def _cifar(store, name, n=128, classes=10):
    rng = np.random.default_rng(0)
    store.create(name, rng.integers(0, 255, (n, 32, 32, 3), dtype=np.uint8), rng.integers(0, classes, (n,)),
                 rng.integers(0, 255, (64, 32, 32, 3), dtype=np.uint8), rng.integers(0, classes, (64,)))


def _run(cluster, req):
    from kubeml_b200.controller import make_kubeml_client
    jid = make_kubeml_client(cluster).v1().networks().train(req)
    return cluster.wait(jid, 900), cluster.jobs[jid]


def test_resnet32_function(cluster):
    _cifar(cluster.shard_store, "cifar10")
    req = TrainRequest(model_type="resnet32", batch_size=32, epochs=1, dataset="cifar10", lr=0.05,
                       function_name="resnet32",
                       options=TrainOptions(default_parallelism=2, static_parallelism=True, k=1))
    err, job = _run(cluster, req)
    assert err is None and np.isfinite(job.history.train_loss[0]) and len(job.history.accuracy) == 1


def test_vgg_function_with_adam(cluster, monkeypatch):
    monkeypatch.setenv("KUBEML_VGG_ARCH", "vgg11")        # same code path, a fraction of the CPU time
    monkeypatch.setenv("KUBEML_VGG_CLASSES", "100")
    _cifar(cluster.shard_store, "cifar100", n=64, classes=100)
    req = TrainRequest(model_type="vgg", batch_size=32, epochs=1, dataset="cifar100", lr=1e-4,
                       function_name="vgg16",
                       options=TrainOptions(default_parallelism=1, static_parallelism=True, k=-1))
    err, job = _run(cluster, req)
    assert err is None and np.isfinite(job.history.train_loss[0])


def test_resnet34_function_infer(cluster, monkeypatch):
    monkeypatch.setenv("KUBEML_RESNET_CLASSES", "10")
    monkeypatch.setenv("KUBEML_RESNET_IMPL", "compat")
    _cifar(cluster.shard_store, "cifar10", n=64)
    from kubeml_b200.runtime.services import Services
    from kubeml_b200.runtime.worker import Worker
    w = Worker(Services(shard_store=cluster.shard_store, tensor_store=cluster.tensor_store))
    q = {"task": "init", "jobId": "j", "N": 1, "K": -1, "funcId": 0, "batchSize": 8, "lr": 0.1, "epoch": 1}
    layers, status = w.invoke("resnet34", q)
    assert status == 200 and "fc.weight" in layers and cluster.tensor_store.exists("j:fc.weight")
    preds, status = w.invoke("resnet34", dict(q, task="infer"), body=np.zeros((2, 32, 32, 3), np.uint8).tolist())
    assert status == 200 and len(preds["predictions"]) == 2
    bad, status = w.invoke("resnet34", dict(q, task="infer"), body=None)
    assert status == 400 and bad["code"] == 400
    unk, status = w.invoke("resnet34", dict(q, task="nonsense"))
    assert status == 400
