#!/usr/bin/env python
"""The pull-fused GEMM alone for Nsight Compute, one GPU (the owner of every weight tile is this GPU, so the report
shows the kernel's pipeline -- tensor pipe, TMA, stalls -- not NVLink): forward with four M-tiles per CTA, the data
gradient with four M-tiles x split-K 8, and the one-M-tile forward of a small batch.

    python tools/profile_pulled.py                      # must exit 0 first
    ncu --set full --import-source on --clock-control none -k regex:umma_pulled -c 3 -f -o gpurun_out/pulled_v2 \
        python tools/profile_pulled.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from mpit_b200.ops.pulled_linear import PulledLinear
from mpit_b200.parallel.symm import SymmetricHeap

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
heap = SymmetricHeap(dev)
lin = PulledLinear(heap, 1024, 4096, 512, act=1)
lin.load_rows(torch.randn(4096, 1024, device=dev) * 0.02)
x, dy = torch.randn(512, 1024, device=dev), torch.randn(512, 4096, device=dev) * 0.1
bias = torch.randn(4096, device=dev) * 0.1
y = lin.forward(x, bias)                  # 64 CTAs x 4 M-tiles, weight ring 11 deep
dx = lin.backward(dy)                     # 16 output tiles x split-K 8 = 128 CTAs, red.add epilogue
y1 = lin.forward(x[:128], bias)           # 64 CTAs x 1 M-tile, weight ring 16 deep
torch.cuda.synchronize()
lin.check()
w = lin.weight_rows()
ref = torch.relu(x @ w.t() + bias)
print("forward rel err", float((lin.forward(x, bias) - ref).norm() / ref.norm()),
      "backward rel err", float((dx - dy @ w).norm() / (dy @ w).norm()))
