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
    if os.environ.get("KUBEML_RESNET32_IMPL", "native") != "torch":
        from kubeml_b200.models.cifar_native import resnet32 as native_resnet32
        return native_resnet32(num_classes)
    return resnet32(num_classes).to(memory_format=torch.channels_last)


def main():
    net = build_network(int(os.environ.get("KUBEML_RESNET_CLASSES", "10")))
    ds = Cifar10Dataset(os.environ.get("KUBEML_RESNET_DATASET", "cifar10"))
    ds.native = bool(getattr(net, "native", False))
    return KubeResnet32(net, ds, gpu=torch.cuda.is_available()).start()
