"""All-reduce tuning sweep: two-shot time vs number of CTAs (torchrun, N GPUs)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
from mpit_b200.parallel import SymmetricHeap, DeviceCollectives

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
heap = SymmetricHeap(dev)
flush = torch.empty(160 << 20, dtype=torch.uint8, device=dev)

def timed(fn, iters=5):
    for _ in range(2):
        fn()
    tot = 0.0
    for _ in range(iters):
        flush.fill_(1); torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    t = torch.tensor([tot / iters], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t) * 1e3

for nbytes in (1 << 20, 16 << 20, 256 << 20):
    n = nbytes // 4
    coll = DeviceCollectives(heap, n)
    x = torch.randn(n, device=dev)
    buf = torch.randn(n, device=dev)
    us_nccl = timed(lambda: dist.all_reduce(buf))
    row = {"bytes": nbytes, "nccl_us": round(us_nccl, 1)}
    for blocks in (16, 32, 64, 148, 296):
        row[f"two_{blocks}"] = round(timed(lambda: coll.allreduce(n=n, algo="twoshot", max_blocks=blocks)), 1)
    for blocks in (32, 148, 296):
        row[f"ela_{blocks}"] = round(timed(lambda: coll.allreduce(n=n, algo="twoshot", average=True, elastic_x=x, alpha=0.1, max_blocks=blocks)), 1)
    if nbytes <= (16 << 20):
        for blocks in (32, 148):
            row[f"one_{blocks}"] = round(timed(lambda: coll.allreduce(n=n, algo="oneshot", max_blocks=blocks)), 1)
    coll.check_timeout()
    if rank == 0:
        print(json.dumps(row), flush=True)
    del coll
dist.barrier(); dist.destroy_process_group()
