"""CIFAR ResNet-32 on CIFAR-10-shaped data: the reference's `resnet32` experiment
(ml/experiments/kubeml/resnet32.py + the function_resnet34 recipe with SGD momentum 0.9)."""
from __future__ import annotations

import os

import torch

from kubeml_b200.functions.function_resnet34 import Cifar10Dataset, KubeResnet34
from kubeml_b200.models.cifar_resnet import resnet32
from kubeml_b200.optim import SGD


class KubeResnet32(KubeResnet34):
    def configure_optimizers(self):
        return SGD(self.parameters(), lr=self.lr, momentum=0.9, weight_decay=1e-4)


def build_network(num_classes: int = 10):
    """The hand-written-kernel model by default; KUBEML_RESNET32_IMPL=torch selects the stock modules
    (same parameter names, so checkpoints and merges are interchangeable)."""
    depth = os.environ.get("KUBEML_CIFAR_RESNET", "resnet32")      # resnet20/32/44/56/110/1202 (resnet32.py:209-234)
    if depth not in ("resnet20", "resnet32", "resnet44", "resnet56", "resnet110", "resnet1202"):
        raise ValueError(f"KUBEML_CIFAR_RESNET={depth}: not one of the CIFAR ResNets")
    if os.environ.get("KUBEML_RESNET32_IMPL", "native") != "torch":
        from kubeml_b200.models import cifar_native
        return getattr(cifar_native, depth)(num_classes)
    from kubeml_b200.models import cifar_resnet
    return getattr(cifar_resnet, depth)(num_classes).to(memory_format=torch.channels_last)


def main():
    net = build_network(int(os.environ.get("KUBEML_RESNET_CLASSES", "10")))
    ds = Cifar10Dataset(os.environ.get("KUBEML_RESNET_DATASET", "cifar10"))
    ds.native = bool(getattr(net, "native", False))
    return KubeResnet32(net, ds, gpu=torch.cuda.is_available()).start()
