"""Native model zoo: the reference's example networks (ml/experiments/kubeml/) rebuilt on the
hand-written sm_100a kernels.  `resnet34` keeps torchvision's parameter names and shapes."""
from .resnet import NativeResNet, resnet18, resnet34

__all__ = ["NativeResNet", "resnet18", "resnet34"]
