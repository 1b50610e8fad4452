import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from mpit_b200.models.cifar_cnn import *
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
torch.manual_seed(4)
model = TorchCifarCNN()
theta = flat_from_torch(model).cuda()
images = torch.randint(0, 256, (B, 3, 32, 32), dtype=torch.uint8)
labels = torch.randint(0, NUM_CLASSES, (B,), dtype=torch.int32)
crop = random_crop_offsets(B)
if os.environ.get('CROP0'): crop = crop * 0 + int(os.environ['CROP0']) - 1
ref = CifarEngine(B, impl="ref", seed=3)
um = CifarEngine(B, impl="umma", seed=3)
for e in (ref, um):
    e.load_flat(theta.clone()); e.set_batch(images, labels, crop)
ref.fwd_bwd()
um.ops.prep_u8(um.x_u8, um.x0, B, 32, 32)
torch.cuda.synchronize()
for i in range(5):
    um.act[i].copy_(ref.act[i]); um.dz[i].copy_(ref.dz[i])
um.gpk.fill_(7.0)
um._umma.wgrad.part.fill_(5.0)
for i in range(4, -1, -1):
    um._umma.wgrad.run(i)
torch.cuda.synchronize()
print("err flag", int(um._umma.err.item()))
rl = lambda a, b: float((a - b).norm() / (b.norm() + 1e-12))
for i in range(5):
    L = um.layers[i]
    gw_u = um.gw(i).view(L.CO, L.k, L.KWP, L.C)[:, :, :L.k]
    gw_r = ref.gw(i).view(L.CO, L.k, L.KWP, L.C)[:, :, :L.k]
    l = um._umma.wgrad.launches[i]
    part = um._umma.wgrad.part[l.part_off:l.part_off + l.nsplit * l.m_pad * l.N].view(l.nsplit, l.m_pad, l.N)
    print(i, "gw rel", rl(gw_u, gw_r), "gb rel", rl(um.gb(i), ref.gb(i)), "|gw_u|", float(gw_u.norm()), "|gw_r|", float(gw_r.norm()),
          "part nz frac", float((part != 0).float().mean()), "part==5 frac", float((part == 5.0).float().mean()),
          "part absmax", float(part.abs().max()))
    # ratio check
    m = gw_r.abs() > gw_r.abs().max() * 0.1
    if m.any():
        print("   ratio sample", (gw_u[m] / gw_r[m])[:6].tolist())
    if i == 0:
        for kh in range(5):
            print("   kh", kh, "rel", rl(gw_u[:, kh], gw_r[:, kh]))
        for kw in range(5):
            print("   kw", kw, "rel", rl(gw_u[:, :, kw], gw_r[:, :, kw]))
        # is it a kw shift?
        for sh in (-2, -1, 1, 2):
            a = gw_u[:, :, max(0, sh):5 + min(0, sh)]; b = gw_r[:, :, max(0, -sh):5 + min(0, -sh)]
            print("   shift", sh, rl(a, b))
