"""Native model zoo: the reference's example networks (ml/experiments/kubeml/) rebuilt on the
hand-written sm_100a kernels.  All keep the reference's / torchvision's parameter names and shapes."""
from . import cifar_native
from .cifar_native import NativeCifarResNet
from .lenet import NativeLeNet, lenet
from .resnet import NativeResNet, resnet18, resnet34
from .vgg import NativeVGG, vgg11, vgg13, vgg16, vgg19

__all__ = ["cifar_native", "NativeCifarResNet", "NativeLeNet", "lenet", "NativeResNet", "resnet18", "resnet34", "NativeVGG", "vgg11", "vgg13", "vgg16",
           "vgg19"]
