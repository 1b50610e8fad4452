"""test: a float passed once around the ring of ranks (reference /root/reference/test.lua:1-42: rank 0 sends
first, everybody else receives then forwards; the only check there is ``assert(size>1)`` :17 -- here the value that
arrives is checked too).

    python -m mpit_b200.launch.run -np 2 -m mpit_b200.launch.test                  # host message layer (MPI_Send/Recv)
    python -m mpit_b200.launch.run -np 2 -m mpit_b200.launch.test --device cuda    # same calls, CUDA tensors, NVLink
"""
from __future__ import annotations

import argparse
import os
import sys

import torch


def _cpu() -> int:
    from .. import msg as mpiT
    mpiT.Init()
    rank, size = mpiT.get_rank(), mpiT.get_size()
    assert size > 1, "run with at least two ranks (test.lua:17)"
    src, dest = (rank - 1) % size, (rank + 1) % size
    s, r = torch.full((1,), float(rank)), torch.full((1,), -1.0)
    st = mpiT.Status()
    print(f"[{rank}/{size}] before: sbuf {float(s[0])} rbuf {float(r[0])}")
    t0 = mpiT.Wtime()
    if rank == 0:
        mpiT.Send(s, 1, mpiT.FLOAT, dest, 0, mpiT.COMM_WORLD)
        mpiT.Recv(r, 1, mpiT.FLOAT, src, 0, mpiT.COMM_WORLD, st)
    else:
        mpiT.Recv(r, 1, mpiT.FLOAT, src, 0, mpiT.COMM_WORLD, st)
        mpiT.Send(s, 1, mpiT.FLOAT, dest, 0, mpiT.COMM_WORLD)
    dt = mpiT.Wtime() - t0
    print(f"[{rank}/{size}] after: rbuf {float(r[0])} from {st.source} ({dt * 1e6:.0f} us)")
    assert float(r[0]) == float(src)
    mpiT.Barrier()
    mpiT.Finalize()
    return 0


def _gpu() -> int:
    """The same script on device buffers: ``mpiT.Send`` / ``mpiT.Recv`` of a CUDA tensor go GPU to GPU through the
    device transport of the message layer (msg/device.py) -- what the reference gets from a CUDA-aware MPI."""
    from .. import msg as mpiT
    from ..msg import device as mdev
    mpiT.Init()
    rank, world = mpiT.get_rank(), mpiT.get_size()
    assert world > 1, "run with at least two ranks (test.lua:17)"
    local = int(os.environ.get("LOCAL_RANK", str(rank % max(torch.cuda.device_count(), 1))))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    mdev.enable()
    src, dest = (rank - 1) % world, (rank + 1) % world
    s, r = torch.full((1,), float(rank), device=dev), torch.full((1,), -1.0, device=dev)
    st = mpiT.Status()
    mpiT.Barrier()
    t0 = mpiT.Wtime()
    if rank == 0:
        mpiT.Send(s, 1, mpiT.FLOAT, dest, 0, mpiT.COMM_WORLD)
        mpiT.Recv(r, 1, mpiT.FLOAT, src, 0, mpiT.COMM_WORLD, st)
    else:
        mpiT.Recv(r, 1, mpiT.FLOAT, src, 0, mpiT.COMM_WORLD, st)
        mpiT.Send(s, 1, mpiT.FLOAT, dest, 0, mpiT.COMM_WORLD)
    dt = mpiT.Wtime() - t0
    print(f"[{rank}/{world}] after: rbuf {float(r[0])} from {st.source} ({dt * 1e6:.0f} us, device transport)")
    assert float(r[0]) == float(src)
    mpiT.Barrier()
    mdev.disable()
    mpiT.Finalize()
    return 0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--device", default="cpu", choices=["cpu", "cuda"])
    a = ap.parse_args(argv)
    return _gpu() if a.device == "cuda" else _cpu()


if __name__ == "__main__":
    sys.exit(main())
