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


def main():
    net = resnet32(int(os.environ.get("KUBEML_RESNET_CLASSES", "10"))).to(memory_format=torch.channels_last)
    ds = Cifar10Dataset(os.environ.get("KUBEML_RESNET_DATASET", "cifar10"))
    return KubeResnet32(net, ds, gpu=torch.cuda.is_available()).start()
