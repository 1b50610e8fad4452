"""ptest: parameter-server bandwidth test (reference /root/reference/asyncsgd/ptest.lua: T rounds of
``async_recv_param + async_send_grad + wait`` over a 671 MB vector, printed as MB/s with a 1-second timer).

* ``--device cpu``  the reference's protocol over the host message layer::

      python -m mpit_b200.launch.run -np 4 -m mpit_b200.launch.ptest --device cpu --ssize 1048576

* ``--device cuda`` the device path: a sweep (default 1 KB .. 1 GB) of the fused exchange kernels --
  pull, push, put, elastic, all-reduce (one-shot / two-shot) -- each timed with CUDA events, max over
  ranks, reported as GB/s per GPU and as a fraction of the measured NVLink peer bandwidth, next to the
  NCCL formulation of the same exchange (``--nccl``)::

      python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m mpit_b200.launch.ptest
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

NVLINK_GBS = 770.0        # measured peer-copy bandwidth per direction per GPU (B200_PROFILING.md)
NVLINK_LINK_GBS = 900.0   # NVLink 5 link bandwidth per direction per GPU: the roofline denominator of BASELINE.json


def _cpu(a, on_gpu: bool = False) -> int:
    """The reference's ptest over the message layer.  ``on_gpu``: every rank takes a GPU, the device transport of the
    message layer is enabled and servers and clients keep their vectors in device memory -- the unmodified
    pServer / pClient protocol (aio_send / aio_recv over Isend / Iprobe / Irecv / Test) then runs GPU to GPU."""
    from .. import msg as mpiT, ps
    mpiT.Init()
    rank, size = mpiT.get_rank(), mpiT.get_size()
    assert size > 0 and size % 2 == 0
    conf = ps.make_conf(rank, size, "halves")
    dev = "cpu"
    if on_gpu:
        from ..msg import device as mdev
        local = rank % torch.cuda.device_count()
        torch.cuda.set_device(local)
        dev = f"cuda:{local}"
        conf["device"] = dev
        mdev.enable()
    if rank < size // 2:
        print(f"rank {rank} is server.")
        ps.pServer(conf).start()
    else:
        print(f"rank {rank} is client.")
        theta, grad = torch.zeros(a.ssize, device=dev), torch.zeros(a.ssize, device=dev)
        pc = ps.pClient(conf)
        pc.start(theta, grad)
        t0 = mpiT.Wtime()
        for t in range(a.rounds):
            pc.async_recv_param()
            pc.async_send_grad()
            pc.wait()
        dt = mpiT.Wtime() - t0
        print(f"rank {rank} bandwidth(bi-direction) is {2 * a.rounds * a.ssize * 4 / dt / 1024 / 1024:.1f} MBytes/sec"
              + (" (device transport)" if on_gpu else ""))
        pc.stop()
        print("pc stopped")
    if on_gpu:
        from ..msg import device as mdev
        st = mdev.transport().stats
        print(f"rank {rank} device transport: {st}")
        mpiT.Barrier()
        mdev.disable()
    mpiT.Finalize()
    return 0


def device_sweep(sizes, iters, warmup, ops, nccl, out_path=None, straggler=0):
    import torch.distributed as dist
    from ..parallel import SymmetricHeap, ShardedCenter, DeviceCollectives
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)
    heap = SymmetricHeap(dev)
    flush = torch.empty(160 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2
    results = []

    def timed(fn):
        """Per-iteration CUDA-event timing with an L2 flush and a rank barrier before each iteration."""
        for _ in range(warmup):
            fn()
        tot = 0.0
        for _ in range(iters):
            flush.fill_(1)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        t = torch.tensor([tot / iters], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t) * 1e-3

    frac = (world - 1) / world if world > 1 else 1.0
    for nbytes in sizes:
        n = max(4, nbytes // 4)
        center = ShardedCenter.create(heap, n)
        x = torch.randn(n, device=dev)
        g = torch.randn(n, device=dev) * 1e-3
        d = torch.empty(n, device=dev)
        coll = DeviceCollectives(heap, n) if any(o.startswith("allreduce") for o in ops) else None
        if rank == 0:
            center.put(x)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        table = {
            "pull": (lambda: center.pull(x), 4 * n * frac),
            "push": (lambda: center.push_add(g), 4 * n * frac),
            "put": (lambda: center.put(x), 4 * n * frac),
            # every rank pulls its remote shards (inbound) AND receives the other ranks' d for the shard it
            # owns (inbound too), and symmetrically outbound: 2 x 4n x frac bytes per direction per GPU
            "elastic": (lambda: center.elastic(x, d, 0.1), 2 * 4 * n * frac),
            "allreduce_oneshot": (lambda: coll.allreduce(n=n, algo="oneshot"), 4 * n * frac),
            "allreduce_twoshot": (lambda: coll.allreduce(n=n, algo="twoshot"), 4 * n * frac),
            "allreduce_elastic": (lambda: coll.allreduce(n=n, average=True, elastic_x=x, alpha=0.1), 4 * n * frac),
        }
        for op in ops:
            if op == "allreduce_oneshot" and n > (8 << 20):
                continue   # latency algorithm: every rank reads everything
            fn, link_bytes = table[op]
            if op == "allreduce_oneshot":
                link_bytes = 4 * n * (world - 1) if world > 1 else 4 * n
            sec = timed(fn)
            row = {"op": op, "bytes": 4 * n, "n_gpus": world, "us": round(sec * 1e6, 2),
                   "link_GBs_per_gpu": round(link_bytes / sec / 1e9, 2),
                   "frac_of_900": round(link_bytes / sec / 1e9 / NVLINK_LINK_GBS, 4) if world > 1 else None,
                   "frac_of_measured_nvlink": round(link_bytes / sec / 1e9 / NVLINK_GBS, 4) if world > 1 else None,
                   "impl": "mpit"}
            results.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
        if nccl and world > 1:
            buf = torch.randn(n, device=dev)
            sec = timed(lambda: dist.all_reduce(buf))
            row = {"op": "allreduce", "bytes": 4 * n, "n_gpus": world, "us": round(sec * 1e6, 2),
                   "link_GBs_per_gpu": round(4 * n * frac / sec / 1e9, 2),
                   "frac_of_900": round(4 * n * frac / sec / 1e9 / NVLINK_LINK_GBS, 4),
                   "frac_of_measured_nvlink": round(4 * n * frac / sec / 1e9 / NVLINK_GBS, 4), "impl": "nccl"}
            results.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
            # the reference's pull as library calls: an all_gather of the shards
            shard = torch.randn((n + world - 1) // world, device=dev)
            full = torch.empty(shard.numel() * world, device=dev)
            sec = timed(lambda: dist.all_gather_into_tensor(full, shard))
            row = {"op": "pull(all_gather)", "bytes": 4 * n, "n_gpus": world, "us": round(sec * 1e6, 2),
                   "link_GBs_per_gpu": round(4 * n * frac / sec / 1e9, 2),
                   "frac_of_900": round(4 * n * frac / sec / 1e9 / NVLINK_LINK_GBS, 4),
                   "frac_of_measured_nvlink": round(4 * n * frac / sec / 1e9 / NVLINK_GBS, 4), "impl": "nccl"}
            results.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
            sec = timed(lambda: dist.reduce_scatter_tensor(shard, full))
            row = {"op": "push(reduce_scatter)", "bytes": 4 * n, "n_gpus": world, "us": round(sec * 1e6, 2),
                   "link_GBs_per_gpu": round(4 * n * frac / sec / 1e9, 2),
                   "frac_of_900": round(4 * n * frac / sec / 1e9 / NVLINK_LINK_GBS, 4),
                   "frac_of_measured_nvlink": round(4 * n * frac / sec / 1e9 / NVLINK_GBS, 4), "impl": "nccl"}
            results.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
        if coll is not None:
            coll.check_timeout()
        del center, coll
    heap.check_timeout()
    if out_path and rank == 0:
        os.makedirs(os.path.dirname(out_path) or ".", exist_ok=True)
        with open(out_path, "w") as f:
            for r in results:
                f.write(json.dumps(r) + "\n")
    return results


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--ssize", type=int, default=10 * 4096 * 4096, help="floats (ptest.lua:3)")
    ap.add_argument("--rounds", type=int, default=100)
    ap.add_argument("--min-bytes", type=int, default=1 << 10)
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--ops", default="pull,push,put,elastic,allreduce_oneshot,allreduce_twoshot,allreduce_elastic")
    ap.add_argument("--nccl", action="store_true", help="also time the NCCL formulation")
    ap.add_argument("--out", default=None)
    ap.add_argument("--protocol", default="kernels", choices=["kernels", "ps"],
                    help="--device cuda: 'kernels' sweeps the fused exchange kernels, 'ps' runs the reference's "
                         "pServer/pClient protocol GPU to GPU over the device transport of the message layer "
                         "(start with python -m mpit_b200.launch.run -np N)")
    a = ap.parse_args(argv)
    if a.device == "cpu":
        return _cpu(a)
    if a.protocol == "ps":
        return _cpu(a, on_gpu=True)
    sizes = []
    b = a.min_bytes
    while b <= a.max_bytes:
        sizes.append(b)
        b *= 4
    if sizes[-1] != a.max_bytes:
        sizes.append(a.max_bytes)
    device_sweep(sizes, a.iters, a.warmup, a.ops.split(","), a.nccl, a.out)
    import torch.distributed as dist
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
