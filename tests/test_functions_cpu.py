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


def test_vgg11_function_of_the_reference(cluster, monkeypatch):
    """ml/experiments/kubeml/function_vgg11.py: VGG-11, CIFAR-10 statistics, no augmentation, Adam."""
    monkeypatch.setenv("KUBEML_VGG_CLASSES", "10")
    _cifar(cluster.shard_store, "cifar10", n=64)
    from kubeml_b200.functions.function_vgg11 import Cifar10Plain
    assert Cifar10Plain.AUGMENT is False and Cifar10Plain.MEAN[0] == 0.4914
    req = TrainRequest(model_type="vgg11", batch_size=32, epochs=1, dataset="cifar10", lr=1e-4,
                       function_name="vgg11",
                       options=TrainOptions(default_parallelism=1, static_parallelism=True, validate_every=1, k=-1))
    err, job = _run(cluster, req)
    assert err is None and np.isfinite(job.history.train_loss[0]) and len(job.history.accuracy) == 1


def test_cifar_resnet_depths_share_parameter_names_with_the_stock_twins(monkeypatch):
    """resnet20 ... resnet1202 (ml/experiments/kubeml/resnet32.py:209-234): native and stock state_dicts line up."""
    from kubeml_b200.models import cifar_native, cifar_resnet
    for depth, blocks in (("resnet20", 3), ("resnet56", 9), ("resnet1202", 200)):
        nat = getattr(cifar_native, depth)(10).state_dict()
        ref = getattr(cifar_resnet, depth)(10).state_dict()
        assert list(nat.keys()) == list(ref.keys())
        assert all(tuple(nat[k].shape) == tuple(ref[k].shape) for k in ref)
        assert f"layer3.{blocks - 1}.conv2.weight" in ref
    from kubeml_b200.functions.function_resnet32 import build_network
    monkeypatch.setenv("KUBEML_CIFAR_RESNET", "resnet44")
    assert "layer1.6.conv1.weight" in build_network(10).state_dict()
    monkeypatch.setenv("KUBEML_CIFAR_RESNET", "resnet33")
    with pytest.raises(ValueError):
        build_network(10)


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


_README_FUNCTION = '''
"""The user function of the reference's README (README.md:120-246), imports adapted, nothing else."""
from typing import Any, List, Tuple, Union

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.optim import SGD

from kubeml_b200 import KubeDataset, KubeModel


class LeNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 6, 5)
        self.conv2 = nn.Conv2d(6, 16, 5)
        self.fc1 = nn.Linear(256, 120)
        self.fc2 = nn.Linear(120, 84)
        self.fc3 = nn.Linear(84, 10)

    def forward(self, x):
        y = F.max_pool2d(F.relu(self.conv1(x)), 2)
        y = F.max_pool2d(F.relu(self.conv2(y)), 2)
        y = y.view(y.shape[0], -1)
        return self.fc3(F.relu(self.fc2(F.relu(self.fc1(y)))))


class MnistDataset(KubeDataset):
    def __init__(self):
        super().__init__("mnist")

    def __getitem__(self, index):
        x = torch.as_tensor(self.data[index], dtype=torch.float32).div(255.0).sub(0.1307).div(0.3081)
        return x.view(1, 28, 28), int(self.labels[index])

    def __len__(self):
        return len(self.data)


class KubeLeNet(KubeModel):
    def __init__(self, network: nn.Module, dataset: MnistDataset):
        super().__init__(network, dataset, gpu=False)

    def configure_optimizers(self) -> torch.optim.Optimizer:
        sgd = SGD(self.parameters(), lr=self.lr, weight_decay=1e-4)
        return sgd

    def init(self, model: nn.Module):
        def init_weights(m: nn.Module):
            if isinstance(m, nn.Conv2d):
                nn.init.xavier_uniform_(m.weight)
                nn.init.constant_(m.bias, 0.01)
            if isinstance(m, nn.Linear):
                nn.init.xavier_uniform_(m.weight)
                nn.init.constant_(m.bias, 0.01)

        model.apply(init_weights)

    def train(self, batch, batch_index) -> float:
        x, y = batch
        self.optimizer.zero_grad()
        output = self(x)
        loss = F.cross_entropy(output, y)
        loss.backward()
        self.optimizer.step()
        if batch_index % 10 == 0:
            self.logger.info(f"Index {batch_index}, error: {loss.item()}")
        return loss.item()

    def validate(self, batch, batch_index) -> Tuple[float, float]:
        x, y = batch
        output = self(x)
        loss = F.cross_entropy(output, y)
        pred = output.argmax(dim=1, keepdim=True)
        correct = pred.eq(y.view_as(pred)).sum().item()
        accuracy = 100. * correct / self.batch_size
        self.logger.debug(f'accuracy {accuracy}')
        return accuracy, loss.item()

    def infer(self, data: List[Any]) -> Union[torch.Tensor, np.ndarray, List[float]]:
        pass


def main():
    lenet = LeNet()
    dataset = MnistDataset()
    kubenet = KubeLeNet(lenet, dataset)
    return kubenet.start()
'''


def test_reference_readme_function_runs_unmodified(cluster, tmp_path):
    """SURVEY W1 acceptance: the README's user function (per-sample __getitem__ dataset, `init(self, model)`,
    float-returning train/validate, self.logger) deploys with `function create` and trains."""
    code = tmp_path / "readme_lenet.py"
    code.write_text(_README_FUNCTION)
    cluster.functions.create("readme_lenet", str(code))
    rng = np.random.default_rng(0)
    cluster.shard_store.create("mnist", rng.integers(0, 255, (256, 28, 28), dtype=np.uint8), rng.integers(0, 10, (256,)),
                               rng.integers(0, 255, (64, 28, 28), dtype=np.uint8), rng.integers(0, 10, (64,)))
    req = TrainRequest(model_type="lenet", batch_size=32, epochs=2, dataset="mnist", lr=0.05,
                       function_name="readme_lenet",
                       options=TrainOptions(default_parallelism=2, static_parallelism=True, validate_every=1, k=2))
    err, job = _run(cluster, req)
    assert err is None
    h = job.history
    assert len(h.train_loss) == 2 and all(np.isfinite(h.train_loss)) and len(h.accuracy) == 2
    assert abs(h.train_loss[0] - 2.3) < 0.5          # xavier init: ~ln(10) at the start
