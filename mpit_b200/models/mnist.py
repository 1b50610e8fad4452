"""MNIST-shaped models for the plumbing configurations of BASELINE.json.

The reference tree names MNIST in its README and launcher comments
(/root/reference/README.md:57, asyncsgd/claunch.lua:11, goot.lua:17-18) but ships no MNIST model or
loader; these shapes are this framework's own:

* ``MnistMLP``  1x32x32 -> 1024 -> 2048 -> 10 (claunch: CPU clients over the host message layer)
* ``mnist_cnn_layers``  the CIFAR topology with MNIST channel counts for the B200 kernel engine
  (Conv(1->32,5x5)+pool, Conv(32->64,5x5)+pool, Conv(64->64,3x3), Linear(256->256), Linear(256->10))
  on a 28x28 window of a 32x32 image.
"""
from __future__ import annotations

import torch

from .cifar_cnn import torch7_reset


class MnistMLP(torch.nn.Module):
    def __init__(self, hidden: int = 2048):
        super().__init__()
        self.fc1 = torch.nn.Linear(1024, hidden)
        self.fc2 = torch.nn.Linear(hidden, 10)
        torch7_reset(self)

    def forward(self, x):
        return torch.log_softmax(self.fc2(torch.relu(self.fc1(x.flatten(1)))), dim=1)


def synthetic_mnist(n: int, seed: int = 0):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(n, 1, 32, 32, generator=g)
    y = torch.randint(0, 10, (n,), generator=g)
    return x, y


def flat_params(model: torch.nn.Module):
    """``model:getParameters()`` (goot.lua:41): one flat storage that the module's tensors alias."""
    ps = list(model.parameters())
    n = sum(p.numel() for p in ps)
    flat = torch.zeros(n)
    gflat = torch.zeros(n)
    o = 0
    for p in ps:
        k = p.numel()
        flat[o:o + k].copy_(p.detach().reshape(-1))
        p.data = flat[o:o + k].view_as(p)
        p.grad = gflat[o:o + k].view_as(p)
        o += k
    return flat, gflat
