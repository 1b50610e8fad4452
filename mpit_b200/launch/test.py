"""test: a float passed once around the ring of ranks (reference /root/reference/test.lua:1-42: rank 0 sends
first, everybody else receives then forwards; the only check there is ``assert(size>1)`` :17 -- here the value that
arrives is checked too).

    python -m mpit_b200.launch.run -np 2 -m mpit_b200.launch.test                  # host message layer (MPI_Send/Recv)
    python -m torch.distributed.run --nproc-per-node 2 -m mpit_b200.launch.test --device cuda   # NVLink mailbox
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
    import torch.distributed as dist
    from ..parallel import SymmetricHeap, DeviceMailbox
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world > 1, "run with at least two ranks (test.lua:17)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    heap = SymmetricHeap(dev)
    box = DeviceMailbox(heap, 64)
    src, dest = (rank - 1) % world, (rank + 1) % world
    s, r = torch.full((1,), float(rank), device=dev), torch.full((1,), -1.0, device=dev)
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if rank == 0:
        box.send(s, dest, tag=0)
        box.recv(r, src, tag=0)
    else:
        box.recv(r, src, tag=0)
        box.send(s, dest, tag=0)
    e1.record()
    torch.cuda.synchronize()
    print(f"[{rank}/{world}] after: rbuf {float(r[0])} from {src} ({e0.elapsed_time(e1) * 1e3:.0f} us on the device)")
    assert float(r[0]) == float(src)
    heap.check_timeout()
    dist.barrier()
    dist.destroy_process_group()
    return 0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--device", default="cpu", choices=["cpu", "cuda"])
    a = ap.parse_args(argv)
    return _gpu() if a.device == "cuda" else _cpu()


if __name__ == "__main__":
    sys.exit(main())
