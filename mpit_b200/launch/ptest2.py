"""ptest2: endless push/pull with rank-dependent artificial compute skew -- the reference's straggler
simulation (/root/reference/BiCNN/ptest2.lua:57-70: ``(rank+1-size/2)^2`` busy iterations per round, prints
per-round communication time).  Bounded here by ``--rounds``.

    python -m mpit_b200.launch.run -np 4 -m mpit_b200.launch.ptest2 --rounds 5 --ssize 65536
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m mpit_b200.launch.ptest2 --device cuda

``--device cuda``: the same experiment on the device path -- every GPU pushes to and pulls from the asynchronous
NVLink parameter server (parallel/ps_device.py, owners served by their native server threads) with the same
rank-dependent skew; prints per-round exchange time per rank and checks that nothing was lost.
"""
from __future__ import annotations

import argparse
import sys

import torch


def _gpu(a) -> int:
    import os
    import time
    import torch.distributed as dist
    from ..parallel import SymmetricHeap, ShardedCenter
    from ..parallel.ps_device import AsyncDevicePS
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    heap = SymmetricHeap(dev)
    n = a.ssize
    ps = AsyncDevicePS(heap, n, "add", {}, ShardedCenter.create(heap, n))
    if rank == 0:
        ps.center.put(torch.zeros(n, device=dev))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ps.start_server(interval_us=20)
    g = torch.full((n,), 2.0 ** -10, device=dev)
    x = torch.empty(n, device=dev)
    m = torch.randn(a.skew_size * 8, a.skew_size * 8, device=dev)
    skew = int((rank + 1 - world / 2) ** 2)
    for t in range(a.rounds):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ps.push(gflat=g)
        ps.pull(x)
        e1.record()
        for _ in range(skew):                               # ptest2.lua:66-70: slower ranks compute longer
            m = torch.tanh(m @ m * 1e-3)
        torch.cuda.synchronize()
        print(f"rank {rank} round {t} exchange time {e0.elapsed_time(e1) * 1e-3:.6f}s (skew {skew} iterations)")
    if world > 1:
        dist.barrier()
    deadline = time.time() + 20
    while ps.applied < a.rounds * world and time.time() < deadline:
        time.sleep(0.005)
    drains = ps.stop_server()
    ps.pull(x)
    torch.cuda.synchronize()
    want = a.rounds * world * 2.0 ** -10
    ok = float(x.min()) == want and float(x.max()) == want
    print(f"rank {rank}: {ps.applied} updates on my shard, {drains} drains, longest flow-control wait "
          f"{ps.max_wait_us:.1f} us, conservation {'ok' if ok else 'VIOLATED'}")
    ps.check_timeout()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0 if ok else 1


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--ssize", type=int, default=4096 * 4096 * 4)
    ap.add_argument("--rounds", type=int, default=5)
    ap.add_argument("--skew-size", type=int, default=200)
    ap.add_argument("--device", default="cpu", choices=["cpu", "cuda"])
    a = ap.parse_args(argv)
    if a.device == "cuda":
        return _gpu(a)
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
