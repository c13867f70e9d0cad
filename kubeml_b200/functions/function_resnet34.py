"""ResNet-34 on CIFAR-10-shaped data (reference: ml/experiments/kubeml/function_resnet34.py:13-104).

Same recipe: torchvision's `resnet34()` topology exactly as the reference instantiates it (ImageNet
stem, 1000 logits, :101), SGD with weight decay 1e-4 and no momentum (:62), epoch-based LR drop
(:57-60), flip + crop(32, pad 4) + normalise for training, normalise only for validation (:17-30),
accuracy as correct*100/batch_size (:91).

B200-first differences: the model is `kubeml_b200.models.resnet34` (NHWC bf16 activations, fp32
master weights in the arena, hand-written kernels) when KUBEML_RESNET_IMPL=native, else the stock
torchvision module under bf16 autocast; augmentation runs on the GPU on whole batches; loss and
accuracy stay on the device.
"""
from __future__ import annotations

import os
from typing import Tuple

import torch
import torch.nn.functional as F

from kubeml_b200 import KubeDataset, KubeModel
from kubeml_b200 import ops
from kubeml_b200.optim import SGD

_MEAN = (0.485, 0.456, 0.406)
_STD = (0.229, 0.224, 0.225)
_STABLE = {}      # (device, batch size) -> the batch buffer the native path refills in place, outside a worker


def _stable_pool() -> dict:
    """Where the in-place batch buffers live: in the worker's warm state (one worker = one training stream; the
    buffers survive from epoch to epoch with the captured step that reads them, and two jobs whose worker threads
    share a GPU do not share a buffer)."""
    from kubeml_b200.runtime.context import current_invocation, has_invocation
    if has_invocation():
        ws = current_invocation().worker_state
        if ws is not None:
            return ws.setdefault("stable_batches", {})
    return _STABLE


class Cifar10Dataset(KubeDataset):
    MEAN, STD = _MEAN, _STD      # per-channel normalisation of the 0..1 image
    AUGMENT = True               # flip + crop(32, padding 4) on training batches

    def __init__(self, name: str = "cifar10"):
        super().__init__(name)
        self._consts = {}
        self.native = False

    def _norm(self, device):
        c = self._consts.get(device)
        if c is None:
            mean = torch.tensor(self.MEAN, device=device).view(1, 1, 1, 3) * 255.0
            inv = 1.0 / (torch.tensor(self.STD, device=device).view(1, 1, 1, 3) * 255.0)
            c = (mean, inv)
            self._consts[device] = c
        return c

    def transform_batch(self, x: torch.Tensor, y: torch.Tensor):
        """x: uint8 [B,32,32,3] (HWC, as CIFAR is stored).  Native models get normalised bf16 NHWC
        from one fused kernel (flip + crop + normalise); stock modules get the float NCHW view."""
        B = x.shape[0]
        augment = self.AUGMENT and self.is_training() and os.environ.get("KUBEML_RESNET_NOAUG") != "1"   # parity runs: no randomness
        if self.native and x.is_cuda:
            # one output buffer per (device, batch size), refilled in place for every batch on the training
            # stream: the captured step reads it where it is (network.py `_kubeml_stable`), no per-step copy
            pool = _stable_pool()
            out = pool.get((x.device, B))
            if out is None:
                out = pool[(x.device, B)] = torch.empty((B, 32, 32, 3), dtype=torch.bfloat16, device=x.device)
                out._kubeml_stable = os.environ.get("KUBEML_STABLE_BATCH", "1") != "0"
            state = None
            if augment:        # flip + crop(32, padding 4) drawn inside the kernel: one launch, no host-side random tensors
                state = self._consts.get(("aug", x.device))
                if state is None:
                    state = self._consts[("aug", x.device)] = torch.zeros(2, dtype=torch.int64, device=x.device)
            y64 = (y if y.dtype == torch.int64 else y.to(torch.int64)).contiguous()
            if not out._kubeml_stable:
                ops.nn.image_prep(x, out, None, 4, self.MEAN, self.STD, rand_state=state, seed=0x5EED)
                return out, y64
            ybuf = pool.get((x.device, B, "labels"))           # the labels travel in the same launch, same idea
            if ybuf is None:
                ybuf = pool[(x.device, B, "labels")] = torch.empty(B, dtype=torch.int64, device=x.device)
                ybuf._kubeml_stable = True
            ops.nn.image_prep(x, out, None, 4, self.MEAN, self.STD, rand_state=state, seed=0x5EED, labels=y64, labels_out=ybuf)
            return out, ybuf
        if augment:
            # RandomHorizontalFlip + RandomCrop(32, padding=4), vectorised over the batch
            flip = torch.rand(B, device=x.device) < 0.5
            x = torch.where(flip.view(B, 1, 1, 1), x.flip(2), x)
            xp = F.pad(x, (0, 0, 4, 4, 4, 4))
            oh = torch.randint(0, 9, (B,), device=x.device)
            ow = torch.randint(0, 9, (B,), device=x.device)
            ar = torch.arange(32, device=x.device)
            ih = (oh.view(B, 1) + ar).view(B, 32, 1)
            iw = (ow.view(B, 1) + ar).view(B, 1, 32)
            x = xp[torch.arange(B, device=x.device).view(B, 1, 1), ih, iw]
        mean, inv = self._norm(x.device)
        x = (x.to(torch.float32) - mean) * inv
        return x.permute(0, 3, 1, 2), y.to(torch.int64)     # NCHW logical, channels-last physical

    def __getitem__(self, index):
        x, y = self.transform_batch(torch.as_tensor(self.data[index:index + 1]),
                                    torch.as_tensor(self.labels[index:index + 1]))
        return x[0], y[0]

    def __len__(self):
        return len(self.data)


class KubeResnet34(KubeModel):
    graph_safe = True      # train() below never syncs with the host

    def __init__(self, network, dataset: Cifar10Dataset, gpu: bool = True):
        super().__init__(network, dataset, gpu=gpu)
        self._amp = os.environ.get("KUBEML_AMP", "1") != "0"

    def configure_optimizers(self):
        if self.epoch > 80:       # reference schedule (function_resnet34.py:57-60)
            self.lr *= 0.1
        return SGD(self.parameters(), lr=self.lr, weight_decay=1e-4)

    def _forward(self, x):
        if getattr(self._network, "native", False) or x.device.type != "cuda" or not self._amp:
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
        return correct * (100.0 / self.batch_size), loss

    def infer(self, data):
        x = torch.as_tensor(data, dtype=torch.uint8, device=self.device).view(-1, 32, 32, 3)
        self._dataset._eval()
        xb, _ = self._dataset.transform_batch(x, torch.zeros(x.shape[0], dtype=torch.int64, device=x.device))
        return self._forward(xb).argmax(dim=1)


def build_network(num_classes: int = 1000):
    impl = os.environ.get("KUBEML_RESNET_IMPL", "auto")
    if impl in ("native", "auto"):
        try:
            from kubeml_b200.models.resnet import resnet34 as native_resnet34
            return native_resnet34(num_classes=num_classes)
        except ImportError:
            if impl == "native":
                raise
    from torchvision.models.resnet import resnet34
    net = resnet34(num_classes=num_classes)
    return net.to(memory_format=torch.channels_last)


def main():
    net = build_network(int(os.environ.get("KUBEML_RESNET_CLASSES", "1000")))
    init = os.environ.get("KUBEML_RESNET_INIT")         # start from a saved state_dict (parity runs, warm start)
    if init and os.path.isfile(init):
        net.load_state_dict(torch.load(init, map_location="cpu"))
    ds = Cifar10Dataset(os.environ.get("KUBEML_RESNET_DATASET", "cifar10"))
    ds.native = bool(getattr(net, "native", False))
    return KubeResnet34(net, ds, gpu=torch.cuda.is_available()).start()
