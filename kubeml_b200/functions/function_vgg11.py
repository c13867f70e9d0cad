"""VGG-11 on CIFAR-10-shaped data with Adam: the reference's third example function
(ml/experiments/kubeml/function_vgg11.py:16-106), against the KubeModel API as it is today.

The reference file still has the pre-refactoring signatures (`train(self, model, dataset)` running its own
DataLoader loop, :50-75) and cannot run under its own library any more; what it specifies is kept:
torchvision's `vgg11()` topology with 1000 logits (:103), CIFAR-10 statistics for the normalisation (:27-30),
no augmentation, Adam at the request's learning rate (:54), cross-entropy, accuracy = correct / total * 100 (:96).
The per-batch loop, the device placement and the `loss.item()` per batch are the library's job here.
KUBEML_VGG_IMPL=torch selects the stock torchvision module under bf16 autocast instead of the native kernels."""
from __future__ import annotations

import os

import torch

from kubeml_b200.functions.function_resnet34 import Cifar10Dataset
from kubeml_b200.functions.function_vgg16 import KubeVGG, build_network


class Cifar10Plain(Cifar10Dataset):
    MEAN, STD = (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)     # function_vgg11.py:29
    AUGMENT = False                                                    # ToTensor + Normalize only (:27-30)


def main():
    net = build_network("vgg11", int(os.environ.get("KUBEML_VGG_CLASSES", "1000")))
    ds = Cifar10Plain(os.environ.get("KUBEML_VGG_DATASET", "cifar10"))
    ds.native = bool(getattr(net, "native", False))
    return KubeVGG(net, ds, gpu=torch.cuda.is_available()).start()
