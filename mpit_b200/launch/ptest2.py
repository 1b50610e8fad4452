"""ptest2: endless push/pull with rank-dependent artificial compute skew -- the reference's straggler
simulation (/root/reference/BiCNN/ptest2.lua:57-70: ``(rank+1-size/2)^2`` busy iterations per round, prints
per-round communication time).  Bounded here by ``--rounds``.

    python -m mpit_b200.launch.run -np 4 -m mpit_b200.launch.ptest2 --rounds 5 --ssize 65536
"""
from __future__ import annotations

import argparse
import sys

import torch


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--ssize", type=int, default=4096 * 4096 * 4)
    ap.add_argument("--rounds", type=int, default=5)
    ap.add_argument("--skew-size", type=int, default=200)
    a = ap.parse_args(argv)
    from .. import msg as mpiT, ps
    mpiT.Init()
    rank, size = mpiT.get_rank(), mpiT.get_size()
    conf = ps.make_conf(rank, size, "halves")
    if rank < size // 2:
        ps.pServer(conf).start()
    else:
        theta, grad = torch.zeros(a.ssize), torch.zeros(a.ssize)
        pc = ps.pClient(conf)
        pc.start(theta, grad)
        m = torch.randn(a.skew_size, a.skew_size)
        times = []
        for t in range(a.rounds):
            t0 = mpiT.Wtime()
            pc.async_recv_param()
            pc.async_send_grad()
            pc.wait()
            times.append(mpiT.Wtime() - t0)
            for _ in range((rank + 1 - size // 2) ** 2):    # ptest2.lua:66-70: slower ranks compute longer
                m = torch.tanh(m @ m * 1e-3)
            print(f"rank {rank} round {t} comm time {times[-1]:.4f}s")
        pc.stop()
    mpiT.Finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
