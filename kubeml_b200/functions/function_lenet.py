"""LeNet on MNIST-shaped data (reference: ml/experiments/kubeml/function_lenet.py:14-139).

Same network (1->6->16 conv5, fc 256->120->84->10, ReLU after the last fc as the reference has
it, :36-50), same optimiser (SGD momentum 0.9, wd 1e-4, :79-81), same accuracy definition
(correct*100/batch_size, :123).  The per-sample torchvision transform (ToTensor + Normalize,
:57-60) is expressed as a vectorised `transform_batch` so it runs on the batch after the async
H2D copy.
"""
from __future__ import annotations

import random
from typing import Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

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
        y = F.relu(self.fc1(y))
        y = F.relu(self.fc2(y))
        return F.relu(self.fc3(y))


class MnistDataset(KubeDataset):
    def __init__(self, name: str = "mnist"):
        super().__init__(name)

    def transform_batch(self, x: torch.Tensor, y: torch.Tensor):
        x = x.to(torch.float32).div_(255.0).sub_(0.1307).div_(0.3081)
        return x.view(x.shape[0], 1, 28, 28), y.to(torch.int64)

    def __getitem__(self, index):
        x, y = self.transform_batch(torch.as_tensor(self.data[index:index + 1]),
                                    torch.as_tensor(self.labels[index:index + 1]))
        return x[0], y[0]

    def __len__(self):
        return len(self.data)


class KubeLeNet(KubeModel):
    graph_safe = True
    def __init__(self, network: nn.Module, dataset: MnistDataset, gpu: bool = False):
        super().__init__(network, dataset, gpu=gpu)

    def configure_optimizers(self) -> torch.optim.Optimizer:
        return torch.optim.SGD(self.parameters(), lr=self.lr, momentum=0.9, weight_decay=1e-4)

    def train(self, batch, batch_index) -> torch.Tensor:
        x, y = batch
        self.optimizer.zero_grad()
        loss = F.cross_entropy(self(x), y)
        loss.backward()
        self.optimizer.step()
        return loss.detach()

    def validate(self, batch, _) -> Tuple[torch.Tensor, torch.Tensor]:
        x, y = batch
        out = self(x)
        loss = F.cross_entropy(out, y)
        correct = out.argmax(dim=1).eq(y).sum()
        return 100.0 * correct / self.batch_size, loss

    def infer(self, data):
        x = torch.as_tensor(data, dtype=torch.float32, device=self.device).view(-1, 1, 28, 28)
        return self(x).argmax(dim=1)


def build_network() -> nn.Module:
    """On a GPU the same network on the hand-written kernels (`kubeml_b200.models.lenet`: identical
    parameter names, shapes and initialisation); `KUBEML_LENET_IMPL=torch` keeps the stock module above,
    which is also what runs on the CPU."""
    import os
    if torch.cuda.is_available() and os.environ.get("KUBEML_LENET_IMPL", "native") != "torch":
        from kubeml_b200.models import lenet
        return lenet()
    return LeNet()


def main():
    torch.manual_seed(42)
    random.seed(42)
    import os
    return KubeLeNet(build_network(), MnistDataset(os.environ.get("KUBEML_LENET_DATASET", "mnist")),
                     gpu=torch.cuda.is_available()).start()
