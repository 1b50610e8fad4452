import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
from mpit_b200.parallel import SymmetricHeap, ShardedCenter
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
heap = SymmetricHeap(dev)
n = 1 << 28
c1 = ShardedCenter.create(heap, n); c2 = ShardedCenter.create(heap, n)
x = torch.randn(n, device=dev); g = torch.randn(n, device=dev) * 1e-3; d = torch.empty(n, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def timed(fn, iters=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
def both():
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s1): c1.pull(x)
    with torch.cuda.stream(s2): c2.push_add(g)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
tp, tq, tb, te = timed(lambda: c1.pull(x)), timed(lambda: c2.push_add(g)), timed(both), timed(lambda: c1.elastic(x, d, 0.1))
if rank == 0: print(f"pull {tp:.3f} ms  push {tq:.3f} ms  pull||push on two streams {tb:.3f} ms  fused elastic {te:.3f} ms")
dist.barrier(); dist.destroy_process_group()
