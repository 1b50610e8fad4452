#!/usr/bin/env python
"""One eager pass over the kernels added in round 2, between cudaProfilerStart/Stop, for Nsight Compute (1 GPU):
the fp32-accurate (3xTF32) CIFAR step, the device-resident BiCNN step with the asynchronous parameter server
(push / poll / apply / pull, inline server), and the pull-fused GEMM with four M-tiles per CTA plus its data gradient.

    ncu --set full --import-source on --clock-control none --profile-from-start off -f -o gpurun_out/round2 \
        python tools/profile_round2.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from mpit_b200.launch.plaunch import parse
from mpit_b200.models.cifar_cnn import random_crop_offsets
from mpit_b200.ops.pulled_linear import PulledLinear
from mpit_b200.parallel.symm import SymmetricHeap
from mpit_b200.workloads import bicnn as W
from mpit_b200.workloads.cifar import CifarOpts, CifarTrainer, synthetic_cifar

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
heap = SymmetricHeap(dev)
B = 128
tr = CifarTrainer(CifarOpts(name="eamsgd", lr=1e-2, mom=0.99, mb=B, su=2, mva=0.9, impl="umma3x", use_graphs=False), dev, heap)
x, y = synthetic_cifar(B, seed=0)
tr.engine.set_batch(x, y, random_crop_offsets(B))
corpus = W.synthetic_corpus(vocab=22354, n_answers=2000, n_train=256, n_eval=16, pool=8, emb_dim=100, seed=0)
opt = parse(["-optimization", "adam", "-batchSize", "4", "-maxnegsample", "100", "-type", "cuda"])
bt = W.BiCNNDeviceTrainer(corpus, opt, dev, 0, heap, server="inline", use_graphs=False)
lin = PulledLinear(heap, 1024, 4096, 512, act=1)
lin.load_rows(torch.randn(4096, 1024, device=dev) * 0.02)
xs, dy = torch.randn(512, 1024, device=dev), torch.randn(512, 4096, device=dev) * 0.1
for _ in range(3):
    tr.device_step()
    bt.train_step([0, 1, 2, 3])
    lin.forward(xs)
    lin.backward(dy)
torch.cuda.synchronize()
torch.cuda.profiler.start()
tr.device_step()
bt.train_step([4, 5, 6, 7])
lin.forward(xs)
lin.backward(dy)
lin.weight_grad()
lin.push(-0.01)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled: cifar 3xTF32 step, bicnn device step, pulled gemm; cifar loss", float(tr.engine.loss_sum.item()) / B)
