"""A/B of programmatic dependent launch on the 1-GPU training step: timing (CUDA events, alternating
rounds inside one process so box-to-box variance cancels) and bit-equality of the trajectory.

    python tools/dev/pdl_ab.py [steps] [rounds]
"""
import json
import sys

import torch

sys.path.insert(0, ".")
from mpit_b200._native import kernels  # noqa: E402
from mpit_b200.parallel.symm import SymmetricHeap  # noqa: E402
from mpit_b200.workloads.cifar import CifarOpts, CifarTrainer  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
B = 128
g = torch.Generator(device="cpu").manual_seed(5)
nb = 352
dx = torch.randint(0, 256, (nb, B, 3, 32, 32), dtype=torch.uint8, generator=g).to(dev)
dy = torch.randint(0, 10, (nb, B), dtype=torch.int32, generator=g).to(dev)
dc = torch.randint(0, 5, (nb, B, 2), dtype=torch.int32, generator=g).to(dev)
heap = SymmetricHeap(dev)


def make(pdl):
    kernels().mpit_set_pdl(pdl)
    tr = CifarTrainer(CifarOpts(name="eamsgd", lr=1e-2, mom=0.99, mb=B, su=10, mva=0.9, impl="umma",
                                use_graphs=True, seed=1), dev, heap)
    for i in range(35):          # captures every graph variant under this PDL setting
        tr.engine.set_batch(dx[i % nb], dy[i % nb], dc[i % nb])
        tr.device_step()
    torch.cuda.synchronize()
    return tr


def run(tr, n, off=0):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(n):
        j = (off + i) % nb
        tr.engine.set_batch(dx[j], dy[j], dc[j])
        tr.device_step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


trs = {1: make(1), 0: make(0)}
# identical histories so far (35 steps each, same seeds/data): the trajectories must agree bit for bit
p1, p0 = trs[1].params().clone(), trs[0].params().clone()
out = {"bit_equal_after_35": bool(torch.equal(p1, p0)), "max_abs_diff": float((p1 - p0).abs().max())}
times = {0: [], 1: []}
for r in range(rounds):
    for mode in (0, 1):
        times[mode].append(round(run(trs[mode], steps, off=r * steps), 5))
p1, p0 = trs[1].params().clone(), trs[0].params().clone()
out["bit_equal_end"] = bool(torch.equal(p1, p0))
out["finite"] = bool(torch.isfinite(p1).all())
out["ms_per_step_pdl_off"] = times[0]
out["ms_per_step_pdl_on"] = times[1]
heap.check_timeout()
print(json.dumps(out))
