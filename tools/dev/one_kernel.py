"""Time one launch object of the CIFAR plan as a CUDA-graph replay: python tools/dev/one_kernel.py dgrad 1"""
import sys
sys.path.insert(0, ".")
import torch
from mpit_b200.models.cifar_cnn import *
kind, i = sys.argv[1], int(sys.argv[2])
B = 128
eng = CifarEngine(B, impl="umma", seed=3)
eng.load_flat(flat_from_torch(TorchCifarCNN()).cuda())
eng.set_batch(torch.randint(0, 256, (B, 3, 32, 32), dtype=torch.uint8), torch.randint(0, 10, (B,), dtype=torch.int32), random_crop_offsets(B))
for _ in range(3):
    eng.fwd_bwd()
torch.cuda.synchronize()
pl = eng._umma
ref = eng.dz[i - 1].clone() if kind == "dgrad" else None
fn = {"dgrad": lambda: pl.dgrad[i].run(pl.err), "fwd": lambda: pl.fwd[i].run(pl.err), "wgrad": lambda: pl.wgrad.run(i)}[kind]
fn(); torch.cuda.synchronize(); pl.check()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    fn()
for _ in range(20): g.replay()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(300): g.replay()
e1.record(); torch.cuda.synchronize()
chk = float(eng.dz[i - 1].double().sum()) if kind == "dgrad" else float(eng.act[i].double().sum()) if kind == "fwd" else 0.0
print(kind, i, f"{e0.elapsed_time(e1) / 300 * 1e3:.2f} us  checksum {chk:.6e}")
