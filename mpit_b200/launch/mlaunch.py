"""mlaunch: distributed EAMSGD / EASGD / Downpour training (reference /root/reference/asyncsgd/mlaunch.lua:
even ranks = servers, odd ranks = clients, ``eamsgd, su=2, mva=0.9/p, lr=1e-2, mom=0.99``).

Two paths:

* ``--device cuda`` (default when GPUs are present): one process per B200, every GPU is a worker and the
  owner of one shard of the centre (no server processes); start with torchrun::

      python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m mpit_b200.launch.mlaunch

* ``--device cpu``: the reference's process layout over the host message layer::

      python -m mpit_b200.launch.run -np 4 -m mpit_b200.launch.mlaunch --device cpu

``--model mnist`` runs the MNIST-shaped CNN (BASELINE.json config 2), ``--optim downpour`` its Downpour arm,
``--optim sync_easgd`` / ``sync_eamsgd`` the synchronous all-reduce trainer; ``--model pulled_mlp`` trains a
perceptron whose weights are never gathered (pull fused into the GEMMs, workloads/pulled_mlp.py).
"""
from __future__ import annotations

import argparse
import os
import sys

import torch


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--optim", default="eamsgd", choices=["eamsgd", "easgd", "downpour", "sync_easgd", "sync_eamsgd"])
    ap.add_argument("--model", default="cifar", choices=["cifar", "mnist", "pulled_mlp"])
    ap.add_argument("--su", type=int, default=2)
    ap.add_argument("--lr", type=float, default=None)
    ap.add_argument("--mom", type=float, default=0.99)
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--trsize", type=int, default=2048)
    ap.add_argument("--tesize", type=int, default=512)
    ap.add_argument("--mb", type=int, default=128)
    ap.add_argument("--save", default="results/mlaunch")
    ap.add_argument("--data", default=None, help="directory with the cifar-10-batches-t7 files")
    ap.add_argument("--resume", default=None, help="checkpoint (model-epoch-N.net or flat tensor) to start from")
    ap.add_argument("--tester-last", action="store_true",
                    help="GPU path: the last rank does not train; it is the shadow tester of BiCNN/bicnn.lua:576-597 -- it "
                         "pulls the centre variable from the shard owners, evaluates it and saves it, --max-evals times")
    ap.add_argument("--max-evals", type=int, default=3)
    ap.add_argument("--eval-sleep", type=float, default=0.2)
    ap.add_argument("--center-ckpt", default=None,
                    help="GPU path: file for the centre variable; restored from it when it exists, written to it at the "
                         "end by the shard owners in parallel (one flat Torch7 tensor, readable with --resume)")
    a = ap.parse_args(argv)
    lr = a.lr if a.lr is not None else (1e-4 if a.optim == "downpour" else 1e-2)   # mlaunch.lua:50-60
    if a.device == "cuda":
        return _gpu(a, lr)
    return _cpu(a, lr)


def _gpu(a, lr) -> int:
    import torch.distributed as dist
    from ..parallel.symm import SymmetricHeap
    from ..workloads.cifar import CifarTrainer, CifarOpts, cifar_data, load_checkpoint, train_epochs
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    heap = SymmetricHeap(dev)
    if a.model == "pulled_mlp":
        # weights stay sharded: every GEMM pulls its tiles from the owners in-kernel (workloads/pulled_mlp.py)
        from ..workloads import pulled_mlp
        sync = (lambda: (torch.cuda.synchronize(dev), dist.barrier(), torch.cuda.synchronize(dev))) if world > 1 else None
        pulled_mlp.train(heap, batch=a.mb, steps=max(1, a.epochs) * 200, lr=a.lr if a.lr is not None else 0.05,
                         barrier=sync)
        heap.check_timeout()
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0
    tester = world - 1 if (a.tester_last and world > 1) else -1
    if tester >= 0 and a.optim.startswith("sync_"):
        raise SystemExit("--tester-last needs an asynchronous optimizer (a synchronous all-reduce involves every rank)")
    workers = world - (1 if tester >= 0 else 0)
    opts = CifarOpts(name=a.optim, lr=lr, mom=a.mom, mb=a.mb, su=a.su, mva=0.9 / workers, maxepoch=a.epochs,
                     save=a.save, seed=1 + rank, model=a.model)
    print(f"[client] rank {rank} use gpu {local}")
    tr = CifarTrainer(opts, dev, heap, theta0=load_checkpoint(a.resume) if a.resume else None)
    sync = dist.barrier if world > 1 else None
    if a.center_ckpt and tr.center is not None and os.path.exists(a.center_ckpt):
        tr.center.load(a.center_ckpt, barrier=sync)        # the reference has no way to restore the centre
        if rank == 0:
            print("===> centre restored from", a.center_ckpt)
    ch = 1 if a.model == "mnist" else 3
    train, test, real = cifar_data(a.trsize, a.tesize, seed=10 + rank, channels=ch, root=a.data)
    if rank == 0:
        print("===> data:", "cifar-10-batches-t7" if real else "synthetic")
    if rank == tester:
        # shadow tester: the centre variable -- what EASGD actually returns -- is evaluated while the workers train
        import time
        from .. import t7
        os.makedirs(a.save, exist_ok=True)
        for k in range(a.max_evals):
            theta = tr.center.gather_full()                 # one-sided pull of every shard, no worker is interrupted
            tr.opt.x.copy_(theta)
            res = tr.test(test[0], test[1])
            path = os.path.join(a.save, f"centre-eval-{k}.t7")
            t7.save(path, theta.cpu())                      # flat parameter tensor, as bicnn.lua:590-594 saves it
            print(f"[tester {rank}] eval {k}: centre test loss {res['loss']:.4f} accuracy {res['acc']:.4f} -> {path}")
            time.sleep(a.eval_sleep)
        hist = None
    else:
        hist = train_epochs(tr, train, test, rank=rank, quiet=rank != 0)
    if tester >= 0:
        torch.cuda.synchronize()
        sync()                                             # workers done, tester done
    if a.center_ckpt and tr.center is not None:
        torch.cuda.synchronize()
        if sync:
            sync()                                         # every worker's last push has landed
        tr.center.save(a.center_ckpt, barrier=sync)
    if hist is not None:
        print(f"[rank {rank}] train loss {hist['train_loss'][-1]:.4f} test acc {hist['test_acc'][-1]:.4f} "
          f"{hist['samples_per_s']:.0f} samples/s")
    heap.check_timeout()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def _cpu(a, lr) -> int:
    from .. import msg as mpiT, ps
    from ..models.cifar_cnn import TorchCifarCNN
    from ..models.mnist import MnistMLP, synthetic_mnist
    from ..workloads import goot
    from ..workloads.cifar import synthetic_cifar
    torch.set_num_threads(1)
    mpiT.Init()
    rank, size = mpiT.get_rank(), mpiT.get_size()
    conf = ps.make_conf(rank, size, "parity")
    if rank % 2 == 0:
        print("[server] rank", rank, "use cpu")
        ps.pServer(conf).start()
    else:
        print(f"[client] rank {rank} use cpu")
        p = size // 2
        torch.manual_seed(rank)
        if a.model == "mnist":
            model, crop = MnistMLP(), 0
            tr, te = synthetic_mnist(a.trsize, 10 + rank), synthetic_mnist(a.tesize, 2)
        else:
            model, crop = TorchCifarCNN(), 28
            (x, y), (tx, ty) = synthetic_cifar(a.trsize, 10 + rank, pinned=False), synthetic_cifar(a.tesize, 2, pinned=False)
            tr, te = (x.float() / 255, y), (tx.float() / 255, ty)
        opt = dict(name=a.optim, lr=lr, su=a.su, mva=0.9 / p, mom=a.mom, maxepoch=a.epochs, mb=a.mb, trsize=a.trsize,
                   tesize=a.tesize, save=a.save, seed=rank, crop=crop, quiet=rank != 1)
        hist = goot.run(opt, model, tr, te, pclient=ps.pClient(conf), rank=rank)
        print(f"[client {rank}] train loss {hist['train_loss'][-1]:.4f} test acc {hist['test_acc'][-1]:.4f}")
    mpiT.Finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
