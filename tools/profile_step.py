#!/usr/bin/env python
"""Run a few eager (graph-free) CIFAR EAMSGD steps between cudaProfilerStart/Stop, for ncu.

    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python tools/profile_step.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from mpit_b200.workloads.cifar import CifarTrainer, CifarOpts, synthetic_cifar
from mpit_b200.models.cifar_cnn import random_crop_offsets

B = int(os.environ.get("B", "128"))
steps = int(os.environ.get("STEPS", "2"))
opts = CifarOpts(name="eamsgd", lr=1e-2, mom=0.99, mb=B, su=2, mva=0.9, impl=os.environ.get("MPIT_KERNELS", "umma"),
                 use_graphs=False)
tr = CifarTrainer(opts, torch.device("cuda:0"))
x, y = synthetic_cifar(B, seed=0)
tr.engine.set_batch(x, y, random_crop_offsets(B))
for _ in range(4):
    tr.device_step()
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(steps):
    tr.device_step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled", steps, "steps; loss", float(tr.engine.loss_sum.item()) / B)
