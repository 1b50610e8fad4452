import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from mpit_b200.models.cifar_cnn import *
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
torch.manual_seed(0)
theta = flat_from_torch(TorchCifarCNN()).cuda()
images = torch.randint(0, 256, (B, 3, 32, 32), dtype=torch.uint8)
labels = torch.randint(0, 10, (B,), dtype=torch.int32)
crop = random_crop_offsets(B)
engs = {}
for impl in ("ref", "umma"):
    e = CifarEngine(B, impl=impl, seed=7)
    e.load_flat(theta)
    e.set_batch(images, labels, crop)
    e.fwd_bwd()
    torch.cuda.synchronize()
    engs[impl] = e
r, u = engs["ref"], engs["umma"]
for i in range(5):
    db = (r.gb(i) - u.gb(i)).abs().max().item()
    dw = (r.gw(i) - u.gw(i)).abs().max().item()
    print(f"layer {i}: |gb| ref {r.gb(i).abs().max().item():.4e} diff {db:.3e}   |gw| ref {r.gw(i).abs().max().item():.4e} diff {dw:.3e}")
    if db > 1e-3:
        print("   ref", r.gb(i)[:8].tolist()); print("   umma", u.gb(i)[:8].tolist())
        cs = (u.dz[i].view(-1, u.layers[i].CO).sum(0) - u.gb(i)).abs().max().item()
        print("   umma colsum(dz) - gb:", cs)
for li, (cp, tiles, ncols, C) in sorted(u._umma.bias_parts.items()):
    v = cp.view(tiles, ncols)
    print("layer", li, "tiles", tiles, "ncols", ncols, "C", C, "colpart absmax", v.abs().max().item(), "folded sum[:4]",
          v.sum(0).view(-1, C).sum(0)[:4].tolist(), "dz colsum[:4]", u.dz[li].view(-1, C).sum(0)[:4].tolist())
print("nred", u._umma.wgrad._nred, "bias_from_epilogue", u._umma.wgrad.bias_from_epilogue)
