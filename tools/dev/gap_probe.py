"""How much of a device-timed step is the input staging between graph launches?"""
import sys, json
import torch
sys.path.insert(0, ".")
from mpit_b200.parallel.symm import SymmetricHeap
from mpit_b200.workloads.cifar import CifarOpts, CifarTrainer
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
B, nb = 128, 352
dx = torch.randint(0, 256, (nb, B, 3, 32, 32), dtype=torch.uint8, device=dev)
dy = torch.randint(0, 10, (nb, B), dtype=torch.int32, device=dev)
dc = torch.randint(0, 5, (nb, B, 2), dtype=torch.int32, device=dev)
tr = CifarTrainer(CifarOpts(name="eamsgd", lr=1e-2, mom=0.99, mb=B, su=10, mva=0.9), dev, SymmetricHeap(dev))
def run(n, stage):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for i in range(n):
        if stage == 3:
            tr.engine.set_batch(dx[i % nb], dy[i % nb], dc[i % nb])
        elif stage == 1:
            tr.engine.x_u8.copy_(dx[i % nb])
        tr.device_step()
    e1.record(); torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) / n, 5)
for i in range(40): run(1, 3)
out = {}
for rep in range(2):
    for stage in (0, 1, 3):
        out.setdefault(f"copies={stage}", []).append(run(300, stage))
print(json.dumps(out))
