#!/usr/bin/env python
"""Point-to-point numbers of the message API on device tensors (the device transport of the message layer,
mpit_b200/msg/device.py): half round-trip time of a Send / Recv ping-pong and one-way bandwidth of back-to-back sends,
by message size -- eager slots up to 64 KiB, CUDA-IPC rendezvous above.  Host wall clock around blocking calls, as an
MPI user would measure it (the reference's ptest does the same, asyncsgd/ptest.lua:40-55).

    python -m mpit_b200.launch.run -np 2 tools/bench_msg_device.py [out.jsonl]
    python -m mpit_b200.launch.run -np 2 tools/bench_msg_device.py --cpu [out.jsonl]    # same calls, host tensors
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from mpit_b200 import msg as mpiT
from mpit_b200.msg import device as mdev

args = [x for x in sys.argv[1:] if x != "--cpu"]
on_gpu = "--cpu" not in sys.argv[1:]
mpiT.Init()
rank, world = mpiT.get_rank(), mpiT.get_size()
assert world == 2, "run with two ranks"
if on_gpu:
    local = rank % max(torch.cuda.device_count(), 1)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    mdev.enable()
else:
    dev = torch.device("cpu")
sync = torch.cuda.synchronize if on_gpu else (lambda: None)
peer = 1 - rank
st = mpiT.Status()
rows = []
for nbytes in (4, 4096, 64 << 10, 1 << 20, 16 << 20, 256 << 20 if on_gpu else 64 << 20):
    n = nbytes // 4
    a, b = torch.full((n,), float(rank + 1), device=dev), torch.zeros(n, device=dev)
    iters = 200 if nbytes <= (64 << 10) else (50 if nbytes <= (16 << 20) else 10)

    def pingpong(k):
        for _ in range(k):
            if rank == 0:
                mpiT.Send(a, n, mpiT.FLOAT, peer, 1, mpiT.COMM_WORLD)
                mpiT.Recv(b, n, mpiT.FLOAT, peer, 1, mpiT.COMM_WORLD, st)
            else:
                mpiT.Recv(b, n, mpiT.FLOAT, peer, 1, mpiT.COMM_WORLD, st)
                mpiT.Send(a, n, mpiT.FLOAT, peer, 1, mpiT.COMM_WORLD)

    def stream(k):                                   # rank 0 sends k messages back to back, rank 1 acknowledges the last
        reqs = mpiT.Request(k)
        for i in range(k):
            if rank == 0:
                mpiT.Isend(a, n, mpiT.FLOAT, peer, 2, mpiT.COMM_WORLD, reqs, i)
            else:
                mpiT.Irecv(b, n, mpiT.FLOAT, peer, 2, mpiT.COMM_WORLD, reqs, i)
        mpiT.Waitall(reqs)
        if rank == 0:
            mpiT.Recv(b, 1, mpiT.FLOAT, peer, 3, mpiT.COMM_WORLD, st)
        else:
            sync()
            mpiT.Send(a, 1, mpiT.FLOAT, peer, 3, mpiT.COMM_WORLD)

    pingpong(5)
    sync()
    mpiT.Barrier()
    t0 = mpiT.Wtime()
    pingpong(iters)
    sync()
    half_rtt = (mpiT.Wtime() - t0) / iters / 2
    assert float(b[0]) == float(peer + 1) and float(b[-1]) == float(peer + 1)
    stream(3)
    mpiT.Barrier()
    k = max(4, iters // 4)
    t0 = mpiT.Wtime()
    stream(k)
    dt = mpiT.Wtime() - t0
    row = {"bytes": nbytes, "half_round_trip_us": round(half_rtt * 1e6, 1), "one_way_GBps": round(k * nbytes / dt / 1e9, 2),
           "messages": k}
    rows.append(row)
    if rank == 0:
        print(json.dumps(row), flush=True)
if rank == 0:
    print("device transport:", mdev.transport().stats if on_gpu else "off (host tensors)", flush=True)
    if args:
        with open(args[0], "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
mpiT.Barrier()
if on_gpu:
    mdev.disable()
mpiT.Finalize()
