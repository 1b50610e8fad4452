"""A model trained THROUGH the pull-fused GEMMs: an MNIST-shaped two-layer perceptron (1024 -> hidden -> 10) whose two
weight matrices live row-sharded in the GPUs' symmetric memory and are never gathered.  Every forward GEMM fetches its
weight tiles from their owners inside the kernel (``ops.pulled_linear.PulledLinear.forward``), every data gradient
reduces across the owners' transposed shards (``backward``), and every worker adds ``-lr * dW`` into the owners' rows
with ``red.add`` (``push``) without a lock -- the Downpour / hogwild scheme of
``/root/reference/asyncsgd/optim-downpour.lua:46-54`` + ``pserver.lua:83-98`` with the parameter pull
(``pclient.lua:74-84``) folded into the consuming GEMM (``asyncsgd/models/7-layers-cnn.lua:20`` is the reference's
first Linear).  The loss (log-softmax + NLL) is the framework's fused kernel on the 10 live logits.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m mpit_b200.launch.mlaunch --model pulled_mlp
"""
from __future__ import annotations

import time

import torch

from ..ops.pulled_linear import PulledLinear


class PulledMLP:
    def __init__(self, heap, batch: int, in_features: int = 1024, hidden: int = 2048, classes: int = 10, seed: int = 1):
        S = heap.nranks
        self.S, self.rank, self.B, self.classes = S, heap.rank, batch, classes
        self.out_pad = 64 * S                                   # logits padded so that every owner has whole tiles
        self.l1 = PulledLinear(heap, in_features, hidden, batch, act=1)
        self.l2 = PulledLinear(heap, hidden, self.out_pad, batch, act=0)
        g = torch.Generator().manual_seed(seed)                 # the same initialisation on every rank: owners keep theirs
        w1 = (torch.rand(hidden, in_features, generator=g) * 2 - 1) / in_features ** 0.5
        w2 = torch.zeros(self.out_pad, hidden)
        w2[:classes] = (torch.rand(classes, hidden, generator=g) * 2 - 1) / hidden ** 0.5
        dev = self.l1.x.device
        r1, r2 = hidden // S, self.out_pad // S
        self.l1.load_rows(w1[self.rank * r1:(self.rank + 1) * r1].to(dev))
        self.l2.load_rows(w2[self.rank * r2:(self.rank + 1) * r2].to(dev))
        self.dev = dev

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        self.h = self.l1.forward(x)
        self.logits = self.l2.forward(self.h)
        return self.logits[:, :self.classes]

    def step(self, x: torch.Tensor, y: torch.Tensor, lr: float) -> float:
        """One hogwild step on this worker's mini-batch; returns its mean loss."""
        z = self.forward(x)
        logp = torch.log_softmax(z, dim=1)
        loss = -logp.gather(1, y.long()[:, None]).mean()
        dz = torch.zeros(x.shape[0], self.out_pad, device=self.dev)
        dz[:, :self.classes] = (logp.exp() - torch.nn.functional.one_hot(y.long(), self.classes)) / x.shape[0]
        dh = self.l2.backward(dz) * (self.h > 0).float()         # relu'
        self.l2.weight_grad()
        self.l1.backward(dh)                                     # dx of the first layer (exercised; the input needs none)
        self.l1.weight_grad()
        self.l2.push(-lr)
        self.l1.push(-lr)
        return float(loss)

    @torch.no_grad()
    def accuracy(self, x: torch.Tensor, y: torch.Tensor) -> float:
        ok = 0
        for i in range(0, x.shape[0] - self.B + 1, self.B):
            ok += int((self.forward(x[i:i + self.B]).argmax(1) == y[i:i + self.B].long()).sum())
        return ok / max(1, (x.shape[0] // self.B) * self.B)

    def check(self) -> None:
        self.l1.check()
        self.l2.check()


def train(heap, *, batch: int = 128, steps: int = 200, lr: float = 0.05, hidden: int = 2048, seed: int = 0,
          barrier=None, quiet: bool = False) -> dict:
    """Every rank trains on its own learnable synthetic MNIST-shaped data; the weights are shared through the shards."""
    from .cifar import learnable_cifar
    dev = heap.device
    model = PulledMLP(heap, batch, 1024, hidden)
    if barrier:
        barrier()
    xs, ys = learnable_cifar(16 * batch, seed=100 + seed + heap.rank, pinned=False, channels=1)
    tx, ty = learnable_cifar(8 * batch, seed=7, pinned=False, channels=1)
    feat = lambda t: (t.float().view(t.shape[0], -1) / 255.0 - 0.5).to(dev)
    xs, ys, tx, ty = feat(xs), ys.to(dev), feat(tx), ty.to(dev)
    acc0 = model.accuracy(tx, ty)
    losses = []
    t0 = time.perf_counter()
    for s in range(steps):
        j = (s % 16) * batch
        losses.append(model.step(xs[j:j + batch], ys[j:j + batch], lr / heap.nranks))
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    if barrier:
        barrier()
    acc1 = model.accuracy(tx, ty)
    model.check()
    out = {"loss_first": sum(losses[:10]) / 10, "loss_last": sum(losses[-10:]) / 10, "acc_before": acc0, "acc_after": acc1,
           "steps": steps, "samples_per_s": steps * batch / dt}
    if not quiet:
        print(f"[rank {heap.rank}] pulled MLP: loss {out['loss_first']:.4f} -> {out['loss_last']:.4f}, "
              f"test accuracy {acc0:.3f} -> {acc1:.3f}, {out['samples_per_s']:.0f} samples/s")
    return out
