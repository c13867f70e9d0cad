"""VGG-16 on CIFAR-100-shaped data with Adam.  The reference ships a VGG-11
function written against an older KubeModel API (`train(self, model, dataset)`, which no longer
matches KubeModel.__train; ml/experiments/kubeml/function_vgg11.py:50); this is the same
recipe (torchvision VGG, Adam, CrossEntropy, accuracy = correct/total) against the current API.
The network is `kubeml_b200.models.vgg16` (hand-written kernels, NHWC bf16 activations, weights in
the arena, one-launch Adam with its step count on the device, fused loss, fused K-AVG merge of the
138 M parameters); KUBEML_VGG_IMPL=torch selects the stock torchvision module under bf16 autocast
(the path any user-supplied nn.Module takes)."""
from __future__ import annotations

import os
from typing import Tuple

import torch

from kubeml_b200 import KubeModel, ops
from kubeml_b200.functions.function_resnet34 import Cifar10Dataset
from kubeml_b200.optim import Adam


class Cifar100Dataset(Cifar10Dataset):
    def __init__(self, name: str = "cifar100"):
        super().__init__(name)


class KubeVGG(KubeModel):
    graph_safe = True

    def __init__(self, network, dataset, gpu: bool = True):
        super().__init__(network, dataset, gpu=gpu)

    def configure_optimizers(self):
        return Adam(self.parameters(), lr=self.lr)

    def _forward(self, x):
        if getattr(self._network, "native", False) or x.device.type != "cuda":
            return self(x)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            return self(x)

    def train(self, batch, batch_index):
        x, y = batch
        self.optimizer.zero_grad()
        out = self._forward(x)
        loss = ops.cross_entropy(out if out.dtype == torch.bfloat16 else out.float(), y)
        loss.backward()
        self.optimizer.step()
        return loss.detach()

    def validate(self, batch, batch_index) -> Tuple[torch.Tensor, torch.Tensor]:
        x, y = batch
        out = self._forward(x)
        loss, correct = ops.accuracy_count(out if out.dtype == torch.bfloat16 else out.float(), y)
        return correct * (100.0 / max(int(y.shape[0]), 1)), loss       # correct/total, function_vgg11.py:96

    def infer(self, data):
        x = torch.as_tensor(data, dtype=torch.uint8, device=self.device).view(-1, 32, 32, 3)
        self._dataset._eval()
        xb, _ = self._dataset.transform_batch(x, torch.zeros(x.shape[0], dtype=torch.int64, device=x.device))
        return self._forward(xb).argmax(dim=1)


def build_network(name: str = "vgg16", num_classes: int = 1000):
    if os.environ.get("KUBEML_VGG_IMPL", "native") != "torch":
        from kubeml_b200.models import vgg as native_vgg
        return getattr(native_vgg, name)(num_classes=num_classes)
    import torchvision.models as tvm
    net = getattr(tvm, name)(num_classes=num_classes)
    return net.to(memory_format=torch.channels_last)


def main():
    net = build_network(os.environ.get("KUBEML_VGG_ARCH", "vgg16"), int(os.environ.get("KUBEML_VGG_CLASSES", "1000")))
    ds = Cifar100Dataset(os.environ.get("KUBEML_VGG_DATASET", "cifar100"))
    ds.native = bool(getattr(net, "native", False))
    return KubeVGG(net, ds, gpu=torch.cuda.is_available()).start()
