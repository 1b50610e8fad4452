"""Dataset loaders (reference: /root/reference/asyncsgd/goot.lua:52-84, /root/reference/asyncsgd/cifar10.lua)."""
from .cifar10 import CLASSES, load_cifar10, write_cifar10_t7, find_cifar10

__all__ = ["CLASSES", "load_cifar10", "write_cifar10_t7", "find_cifar10"]
