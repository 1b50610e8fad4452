"""glaunch: single-GPU Nesterov SGD on the CIFAR CNN (reference /root/reference/asyncsgd/glaunch.lua:
``opt.name='sgd', lr=1e-1, mom=0.99``).  This is the minimum end-to-end slice: one model trained on
one B200 with this framework's kernels only.

    python -m mpit_b200.launch.glaunch --epochs 1 --trsize 2048
"""
from __future__ import annotations

import argparse
import sys

import torch


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--trsize", type=int, default=2048)
    ap.add_argument("--tesize", type=int, default=1024)
    ap.add_argument("--mb", type=int, default=128)
    ap.add_argument("--lr", type=float, default=1e-1)
    ap.add_argument("--mom", type=float, default=0.99)
    ap.add_argument("--kernels", default="umma")
    ap.add_argument("--save", default="results/glaunch")
    ap.add_argument("--data", default=None, help="directory with the cifar-10-batches-t7 files (default: search, "
                                                 "else synthetic data)")
    ap.add_argument("--resume", default=None, help="model-epoch-N.net (or a flat parameter tensor) to start from")
    a = ap.parse_args(argv)
    from ..workloads.cifar import CifarTrainer, CifarOpts, cifar_data, load_checkpoint, train_epochs
    opts = CifarOpts(name="sgd", lr=a.lr, mom=a.mom, mb=a.mb, impl=a.kernels, maxepoch=a.epochs, save=a.save)
    print("[client] use gpu", 0)
    tr = CifarTrainer(opts, torch.device("cuda:0"), theta0=load_checkpoint(a.resume) if a.resume else None)
    train, test, real = cifar_data(a.trsize, a.tesize, seed=1, root=a.data)
    print("===> data:", "cifar-10-batches-t7" if real else "synthetic (CIFAR-10 shaped)")
    hist = train_epochs(tr, train, test)
    print(f"final train loss {hist['train_loss'][-1]:.4f}, test acc {hist['test_acc'][-1]:.4f}, "
          f"{hist['samples_per_s']:.0f} samples/s")
    return 0


if __name__ == "__main__":
    sys.exit(main())
