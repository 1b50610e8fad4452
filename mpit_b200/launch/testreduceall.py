"""testreduceall: a short parameter-server exchange followed by an all-reduce on every rank
(reference /root/reference/asyncsgd/testreduceall.lua: ssize = 3*4096^2, 3 rounds, then
``Allreduce(6 floats, SUM)`` on all ranks :72-77), plus the stand-alone collective smoke scripts
(/root/reference/test/testreduceall.lua, test/testireduceall.lua) behind ``--only-collective``.

    python -m mpit_b200.launch.run -np 4 -m mpit_b200.launch.testreduceall --ssize 65536
    python -m torch.distributed.run --nproc-per-node 2 -m mpit_b200.launch.testreduceall --device cuda
"""
from __future__ import annotations

import argparse
import os
import sys

import torch


def _cpu(a) -> int:
    from .. import msg as mpiT, ps
    mpiT.Init()
    rank, size = mpiT.get_rank(), mpiT.get_size()
    if not a.only_collective:
        conf = ps.make_conf(rank, size, "halves")
        if rank < size // 2:
            ps.pServer(conf).start()
        else:
            theta, grad = torch.zeros(a.ssize), torch.zeros(a.ssize)
            pc = ps.pClient(conf)
            pc.start(theta, grad)
            t0 = mpiT.Wtime()
            for _ in range(a.rounds):
                pc.async_recv_param()
                pc.async_send_grad()
                pc.wait()
            dt = mpiT.Wtime() - t0
            print(f"rank {rank} bandwidth(bi-direction) is {2 * a.rounds * a.ssize * 4 / dt / 1024 / 1024:.1f} MBytes/sec")
            pc.stop()
    # every rank: Allreduce of 6 floats (testreduceall.lua:72-77) ...
    x = torch.full((6,), float(rank))
    y = torch.zeros(6)
    mpiT.Allreduce(x, y, 6, mpiT.FLOAT, mpiT.SUM, mpiT.COMM_WORLD)
    assert float(y[0]) == size * (size - 1) / 2
    # ... and the MEGS-sized one of test/testreduceall.lua with the Iallreduce variant
    megs = float(os.environ.get("MEGS", str(a.megs)))
    n = int(megs * 1024 * 1024 / 4)
    torch.manual_seed(123)
    big = torch.rand(n)
    out = torch.zeros(n)
    t0 = mpiT.Wtime()
    mpiT.Allreduce(big, out, n, mpiT.FLOAT, mpiT.SUM)
    t1 = mpiT.Wtime()
    req, st = mpiT.Request(), mpiT.Status()
    out2 = torch.zeros(n)
    mpiT.Iallreduce(big, out2, n, mpiT.FLOAT, mpiT.SUM, mpiT.COMM_WORLD, req)
    completed_early = mpiT.Test(req, st)
    mpiT.Wait(req, st)
    if rank == 0:
        print("num nodes:", size, "allreduce toc:", round(t1 - t0, 4), "a[1]", float(big[0]), "reduced[1]", float(out[0]),
              "iallreduce completed at first Test:", completed_early)
    assert torch.allclose(out, big * size, rtol=1e-6) and torch.equal(out, out2)
    mpiT.Finalize()
    return 0


def _gpu(a) -> int:
    import torch.distributed as dist
    from .ptest import device_sweep
    megs = float(os.environ.get("MEGS", str(a.megs)))
    sizes = [24, int(megs * 1024 * 1024)]
    device_sweep(sizes, 5, 2, ["pull", "push", "allreduce_oneshot", "allreduce_twoshot", "allreduce_elastic"], True)
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--device", default="cpu")
    ap.add_argument("--ssize", type=int, default=3 * 4096 * 4096)
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--megs", type=float, default=10)
    ap.add_argument("--only-collective", action="store_true")
    a = ap.parse_args(argv)
    return _gpu(a) if a.device == "cuda" else _cpu(a)


if __name__ == "__main__":
    sys.exit(main())
