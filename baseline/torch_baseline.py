"""Library-call GPU baseline for bench.py (``--impl torch``): the same model, data path and EAMSGD
algorithm expressed with PyTorch library kernels only -- cuDNN convolutions / cuBLAS GEMMs through
autograd (TF32 allowed, the same precision class as the tcgen05 path), torch elementwise ops for the
Nesterov / elastic math, and NCCL collectives for the exchange (pull = all_gather of the centre shards,
push = reduce_scatter of the elastic differences).  This is "the path that only calls NCCL" that
BASELINE.json names as the baseline; nothing from mpit_b200's kernels or engine is on it.

The reference's own build (Lua/Torch7 + MPI) cannot be installed on this image.
"""
from __future__ import annotations

import json
import time

import torch
import torch.distributed as dist
import torch.nn.functional as F

from mpit_b200.models.cifar_cnn import TorchCifarCNN, flat_from_torch, torch7_reset
from mpit_b200.parallel.center import partition


class TorchEAMSGD:
    def __init__(self, dev, world, rank, B, lr=1e-2, mom=0.99, su=10, mva=0.9, graphs=True, tf32=True):
        torch.backends.cudnn.allow_tf32 = tf32          # False: the libraries' fp32 kernels (the reference's precision)
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.benchmark = True
        self.dev, self.world, self.rank, self.B = dev, world, rank, B
        self.lr, self.mom, self.su, self.mva = lr, mom, su, mva
        g = torch.Generator().manual_seed(1 + rank)
        m = TorchCifarCNN()
        torch7_reset(m, g)
        self.model = m.to(dev).to(memory_format=torch.channels_last)
        self.params = [p for p in self.model.parameters()]
        self.n = sum(p.numel() for p in self.params)
        self.v = [torch.zeros_like(p) for p in self.params]
        self.d = [torch.zeros_like(p) for p in self.params]
        parts = partition(self.n, world)
        self.shard_cap = max(s for _, s in parts)
        self.parts = parts
        self.center_shard = torch.zeros(self.shard_cap, device=dev)
        flat = flat_from_torch(self.model)
        if world > 1:
            dist.broadcast(flat, 0)       # seed: centre := first worker's theta (pclient.lua:131-133)
        off, sz = parts[rank]
        self.center_shard[:sz].copy_(flat[off:off + sz])
        self.t = 0
        self.x_u8 = torch.zeros(B, 3, 32, 32, dtype=torch.uint8, device=dev)
        self.labels = torch.zeros(B, dtype=torch.int64, device=dev)
        self.crop = torch.zeros(B, 2, dtype=torch.int64, device=dev)
        self.loss = torch.zeros((), device=dev)
        self.ar = torch.arange(28, device=dev)
        self.bi = torch.arange(B, device=dev)
        self.graphs = graphs
        self.graph = None
        self.have_d = False

    def set_batch(self, x, y, c):
        self.x_u8.copy_(x, non_blocking=True)
        self.labels.copy_(y, non_blocking=True)
        self.crop.copy_(c, non_blocking=True)

    def _local(self):
        """Nesterov look-ahead, forward/backward, apply (asyncsgd/optim-eamsgd.lua:26-47)."""
        with torch.no_grad():
            torch._foreach_mul_(self.v, self.mom)
            torch._foreach_add_(self.params, self.v)
        rows = self.crop[:, 0:1] + self.ar[None, :]
        cols = self.crop[:, 1:2] + self.ar[None, :]
        x = self.x_u8[self.bi[:, None, None], :, rows[:, :, None], cols[:, None, :]]      # [B,28,28,3]
        x = x.permute(0, 3, 1, 2).float().div_(255.0).contiguous(memory_format=torch.channels_last)
        for p in self.params:
            p.grad = None
        loss = F.nll_loss(self.model(x), self.labels)
        loss.backward()
        self.loss.copy_(loss.detach())
        with torch.no_grad():
            grads = [p.grad for p in self.params]
            torch._foreach_add_(self.params, grads, alpha=-self.lr)
            torch._foreach_add_(self.v, grads, alpha=-self.lr)

    def _sync(self):
        """Elastic exchange with NCCL collectives: pull (all_gather), d = mva (x - centre), push (reduce_scatter)."""
        with torch.no_grad():
            if self.world > 1:
                full = torch.empty(self.shard_cap * self.world, device=self.dev)
                dist.all_gather_into_tensor(full, self.center_shard)
                center = torch.cat([full[r * self.shard_cap: r * self.shard_cap + sz] for r, (_, sz) in enumerate(self.parts)])
            else:
                center = self.center_shard[:self.n]
            x = torch.cat([p.reshape(-1) for p in self.params])
            dflat = (x - center) * self.mva
            if self.world > 1:
                pad = torch.zeros(self.shard_cap * self.world, device=self.dev)
                for r, (off, sz) in enumerate(self.parts):
                    pad[r * self.shard_cap: r * self.shard_cap + sz] = dflat[off:off + sz]
                out = torch.empty(self.shard_cap, device=self.dev)
                dist.reduce_scatter_tensor(out, pad)
                self.center_shard.add_(out)
            else:
                self.center_shard[:self.n].add_(dflat)
            o = 0
            for dd in self.d:
                k = dd.numel()
                dd.copy_(dflat[o:o + k].view_as(dd))
                o += k

    def step(self):
        do_sync = self.t % self.su == 0
        if do_sync:
            self._sync()
        if self.graphs and self.t >= 3:
            if self.graph is None:
                torch.cuda.synchronize()
                self.graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.graph):
                    self._local()
            self.graph.replay()
        else:
            self._local()
        if do_sync:
            with torch.no_grad():
                torch._foreach_sub_(self.params, self.d)     # w -= sug (optim-eamsgd.lua:70)
        self.t += 1


def measure(args, rank, world, dev, blocks: int = 5, tf32: bool = True, e2e: bool = True):
    """Both arms of the library baseline, timed exactly like the product arm in bench.py: `blocks` back-to-back blocks
    of `steps` steps each (median block reported); the end-to-end arm takes its batches from the same native loader
    thread and reads the loss of step t-2 every step, i.e. the same host-side rules as CifarTrainer.train_step."""
    import statistics
    B = args.batch
    from mpit_b200.workloads.cifar import synthetic_cifar
    from mpit_b200.data.loader import BatchLoader
    tr = TorchEAMSGD(dev, world, rank, B, su=args.tau, mva=0.9 / world, graphs=not args.no_graphs, tf32=tf32)
    n_host = 64 * B
    hx, hy = synthetic_cifar(n_host, seed=100 + rank)
    n_dev = 352
    dx = torch.randint(0, 256, (n_dev, B, 3, 32, 32), dtype=torch.uint8, device=dev)
    dy = torch.randint(0, 10, (n_dev, B), dtype=torch.int64, device=dev)
    dc = torch.randint(0, 5, (n_dev, B, 2), dtype=torch.int64, device=dev)
    loader = BatchLoader(hx, hy, B, nslots=4, crop_range=5, seed=7 + rank)
    LAG = 2
    loss_host = torch.zeros(LAG + 1).pin_memory()
    loss_evt = [torch.cuda.Event() for _ in range(LAG + 1)]
    state = {"k": 0, "last": float("nan")}

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def dev_step(i):
        j = i % n_dev
        tr.set_batch(dx[j], dy[j], dc[j])
        tr.step()

    def e2e_step(i):
        x, y, c = loader.next()                      # pinned slot filled by the loader thread
        tr.set_batch(x, y, c)                        # host -> device copy of this step's inputs
        tr.step()
        k = state["k"]
        slot = k % (LAG + 1)
        loss_host[slot:slot + 1].copy_(tr.loss.view(1), non_blocking=True)     # device -> host copy of the loss
        loss_evt[slot].record()
        state["k"] = k + 1
        if k >= LAG:                                 # the host reads the loss of step k-LAG every step
            old = (k - LAG) % (LAG + 1)
            loss_evt[old].synchronize()
            state["last"] = float(loss_host[old])
        return state["last"]

    for i in range(max(args.warmup, 3 * args.tau + 3)):
        dev_step(i)
    barrier()
    dev_ms, e2e_ms = [], []
    for b in range(blocks):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for i in range(args.steps):
            dev_step(b * args.steps + i)
        e1.record()
        barrier()
        dev_ms.append(e0.elapsed_time(e1))
    last = float("nan")
    for i in range(5 if e2e else 0):
        e2e_step(i)
    for b in range(blocks if e2e else 0):
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            last = e2e_step(i)
        barrier()
        e2e_ms.append((time.perf_counter() - t0) * 1e3)
    if not e2e:
        e2e_ms = [float("nan")] * blocks
    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)          # per block: the slowest rank
    dms, ems = statistics.median(t[0].tolist()), statistics.median(t[1].tolist())
    loader.close()
    gb = B * world
    return {"value": gb * args.steps / (dms / 1e3), "e2e": gb * args.steps / (ems / 1e3), "ms_per_step": dms / args.steps,
            "e2e_ms_per_step": ems / args.steps, "blocks_ms": [round(v, 4) for v in t[0].tolist()], "loss": last,
            "h2d_bytes_per_step": int(3 * 32 * 32 * B + 4 * B + 8 * B), "d2h_bytes_per_step": 4}


def run(args, rank, world, dev):
    B = args.batch
    m = measure(args, rank, world, dev)
    gb = B * world
    if rank == 0:
        print(json.dumps({
            "metric": "samples/sec (device-timed, max over ranks) CIFAR-10 CNN EAMSGD tau=10",
            "value": round(m["value"], 1), "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(m["ms_per_step"], 5),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "tf32 (cuDNN/cuBLAS allow_tf32)",
            "data": "synthetic CIFAR-10-shaped uint8 images, random-init weights", "impl": "torch",
            "config": {"model": "CIFAR-10 7-layer CNN", "global_batch": gb, "per_gpu_batch": B, "optimizer": "eamsgd",
                       "tau": args.tau, "parallelism": f"sync-elastic dp{world} via NCCL all_gather/reduce_scatter",
                       "cuda_graphs": not args.no_graphs, "timing": "median of 5 blocks of `steps` steps"},
            "e2e": {"value": round(m["e2e"], 1), "unit": "samples/s", "ms_per_step": round(m["e2e_ms_per_step"], 5),
                    "h2d_bytes_per_step": m["h2d_bytes_per_step"], "d2h_bytes_per_step": 4,
                    "loader": "native worker thread, pinned ring", "d2h": "loss copied every step, read 2 steps late"},
            "gpu_launches": 0, "loss": round(m["loss"], 4)}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0
