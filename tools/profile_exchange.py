#!/usr/bin/env python
"""One process driving TWO GPUs: the exchange kernels of rank 0 (cuda:0) against a centre whose second shard lives
on cuda:1, for Nsight Compute (a single-process program, so ncu's kernel replay works; the all-reduce kernels carry a
cross-rank barrier and cannot be replayed -- they are timed with CUDA events instead, see launch/ptest.py).

    ncu --set full --import-source on --clock-control none -k regex:exchange_kernel -c 8 \
        -o gpurun_out/exchange_n2 python tools/profile_exchange.py

Algorithmic NVLink bytes per call on rank 0 (half of the vector is remote): pull n/2*4 B in; push n/2*4 B out;
put n/2*4 B out; elastic n/2*4 B in + n/2*4 B out.  Compare with nvlrx__bytes / nvltx__bytes of the capture."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from mpit_b200.parallel.center import ShardedCenter, shard_capacity
from mpit_b200.parallel.symm import LocalPeerHeap

n = int(os.environ.get("N", str(16 << 20)))           # 64 MiB vector: 32 MiB remote
bulk = os.environ.get("BULK", "0") == "1"
heap = LocalPeerHeap(2, devices=[0, 1])
cs = [ShardedCenter(b, n) for b in heap.alloc(shard_capacity(n, 2))]
c = cs[0]
c.bulk = bulk and c.bulk
torch.cuda.set_device(0)
x = torch.randn(n, device="cuda:0")
g = torch.randn(n, device="cuda:0") * 1e-3
d = torch.empty(n, device="cuda:0")
c.put(x)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(2):
    c.pull(x)
    c.push_add(g)
    c.put(x)
    c.elastic(x, d, 0.1)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("exchange profile done: n =", n, "bulk =", c.bulk, "remote bytes per direction per call =", 2 * n)
