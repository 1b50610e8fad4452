import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
torch.cuda.init()
dbg = torch.zeros(16 * 2048, dtype=torch.int64, device="cuda")
os.environ["MPIT_IGEMM_DBG"] = str(dbg.data_ptr())
from mpit_b200.models.cifar_cnn import *
B = 128
eng = CifarEngine(B, impl="umma", seed=3)
eng.load_flat(flat_from_torch(TorchCifarCNN()).cuda())
eng.set_batch(torch.randint(0, 256, (B, 3, 32, 32), dtype=torch.uint8), torch.randint(0, 10, (B,), dtype=torch.int32), random_crop_offsets(B))
for _ in range(3):
    eng.fwd_bwd()
torch.cuda.synchronize()
names = ["conv1 fwd", "conv2 fwd", "conv3 fwd", "fc1 fwd", "fc2 fwd"]
def show(name, launch, nct):
    dbg.zero_()
    torch.cuda.synchronize()
    launch.run(eng._umma.err)
    torch.cuda.synchronize()
    d = dbg.view(-1, 16)[:nct].cpu().double()
    t0 = d[:, 0].min()
    rel = (d - d[:, 0:1])
    f = lambda i: f"{rel[:, i].mean() / 1.9:8.0f}"
    print(f"{name:12s} ctas {nct:4d} | ns since CTA start: prologue {f(1)} prodSetup {f(8)} slotFree {f(9)} firstIssued {f(10)} lastTMA {f(2)} firstFull {f(3)} lastMMAissue {f(4)} accReady {f(5)} tmemLd {f(11)} math {f(12)} stores {f(13)} epiDone {f(6)} end {f(7)} | span {(d[:,7].max()-t0)/1.9:8.0f}", flush=True)
cts = [576, 64, 8, 8, 1]
for i in range(5):
    show(names[i], eng._umma.fwd[i], cts[i])
show("fc2 dgrad", eng._umma.dgrad[4], 8)
show("fc1 dgrad", eng._umma.dgrad[3], 8)
show("conv3 dgrad", eng._umma.dgrad[2], 32)
show("conv2 dgrad", eng._umma.dgrad[1], 144)
