import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from mpit_b200.models.cifar_cnn import *
B = 2
torch.manual_seed(4)
um = CifarEngine(B, impl="umma", seed=3)
images = torch.randint(0, 256, (B, 3, 32, 32), dtype=torch.uint8)
# recognizable image: value encodes position: r = x, g = y, b = 7
for y in range(32):
    for x in range(32):
        images[:, 0, y, x] = x; images[:, 1, y, x] = y; images[:, 2, y, x] = 100 + x
labels = torch.zeros(B, dtype=torch.int32)
crop = torch.zeros(B, 2, dtype=torch.int32)
um.set_batch(images, labels, crop)
um.ops.prep_u8(um.x_u8, um.x0, B, 32, 32)
for (oh, ow) in [(5, 6), (5, 7), (2, 1)]:
    for i in range(5): um.dz[i].zero_()
    dz1 = um.dz[0].view(B, 24, 24, 64)
    dz1[0, oh, ow, 0] = 1.0
    um._umma.wgrad.run(0)
    um._umma.wgrad.reduce()
    torch.cuda.synchronize()
    gw = um.gw(0).view(64, 5, 32)[0] * 255.0   # [kh][32 floats]
    x0 = um.x0.view(B + 1, 32, 32, 8)[0] * 255.0
    print("pixel", oh, ow)
    for kh in range(2):
        exp = x0[oh + kh, ow:ow + 4].reshape(-1)
        print("  kh", kh, "got", [round(float(v)) for v in gw[kh]])
        print("       exp", [round(float(v)) for v in exp])
