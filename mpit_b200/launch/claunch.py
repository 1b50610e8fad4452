"""claunch: MNIST-shaped MLP, EASGD/EAMSGD, 1 parameter server + N CPU clients over the host message
layer (BASELINE.json config 1; reference /root/reference/asyncsgd/claunch.lua, which as written cannot
run -- eamsgd without a client asserts, optim-eamsgd.lua:75-76 -- so the roles follow BASELINE.json).

    python -m mpit_b200.launch.run -np 3 -m mpit_b200.launch.claunch [--epochs 1]
"""
from __future__ import annotations

import argparse
import sys

import torch


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--trsize", type=int, default=1024)
    ap.add_argument("--tesize", type=int, default=256)
    ap.add_argument("--mb", type=int, default=128)
    ap.add_argument("--optim", default="eamsgd")
    ap.add_argument("--save", default="results/claunch")
    ap.add_argument("--hidden", type=int, default=2048)
    a = ap.parse_args(argv)
    from .. import msg as mpiT, ps
    from ..models.mnist import MnistMLP, synthetic_mnist
    from ..workloads import goot
    torch.set_num_threads(1)
    mpiT.Init()
    rank, size = mpiT.get_rank(), mpiT.get_size()
    assert size >= 2, "claunch needs one server and at least one client"
    conf = {"rank": rank, "world": mpiT.COMM_WORLD, "sranks": [0], "cranks": list(range(1, size)), "size": size}
    if rank == 0:
        print("[server] rank", rank, "use cpu")
        ps.pServer(conf).start()
    else:
        print(f"[client] rank {rank} use cpu")
        torch.manual_seed(rank)
        model = MnistMLP(a.hidden)
        p = size - 1
        # claunch.lua:10-15: lr=1e-2, mva=lr*1e-4, su=1, mom=0.99
        opt = dict(name=a.optim, lr=1e-2, mva=1e-2 * 1e-4 if a.optim != "easgd" else 0.9 / p, su=1, mom=0.99,
                   maxepoch=a.epochs, mb=a.mb, trsize=a.trsize, tesize=a.tesize, save=a.save, seed=rank, quiet=rank != 1)
        hist = goot.run(opt, model, synthetic_mnist(a.trsize, seed=10 + rank), synthetic_mnist(a.tesize, seed=5),
                        pclient=ps.pClient(conf), rank=rank)
        print(f"[client {rank}] train loss {hist['train_loss'][-1]:.4f} test acc {hist['test_acc'][-1]:.4f}")
    mpiT.Finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
