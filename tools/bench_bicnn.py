#!/usr/bin/env python
"""BASELINE.json config 4: BiCNN answer selection with a server-side rule (parallel Adam by default) at the reference
dimensions (embedding 100, hidden 200, 3000 filters, kernel width 2, V = 22 354, up to 100 candidate negatives per
question; /root/reference/BiCNN/plaunch.lua:46-53, bicnn.lua:30-121), one process per GPU.

    python tools/bench_bicnn.py                                   # 1 GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_bicnn.py --gpus 8

``--impl mpit``  : workloads.bicnn.BiCNNDeviceTrainer -- device-resident step, asynchronous NVLink parameter server.
``--impl torch`` : the NCCL formulation of the same algorithm with library kernels only: torch autograd for the
                   encoder (cuBLAS TF32 GEMMs), all_to_all of the gradient shards, the owner applies one Adam update
                   per client with torch elementwise ops, all_gather of the parameters.  Lock-step by construction.

Prints one JSON line: questions/s over all GPUs, device-timed (CUDA events, max over ranks), plus an end-to-end value
through the public ``train_step`` call (sample indices from pinned host memory every step, one loss read per step).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def make_corpus(W, vocab, answers, train, seed=0):
    return W.synthetic_corpus(vocab=vocab, n_answers=answers, n_train=train, n_eval=32, pool=16, emb_dim=100, seed=seed)


class TorchPS:
    """NCCL formulation (see module docstring)."""

    def __init__(self, corpus, opt, dev, rank, world):
        from mpit_b200.models.bicnn import BiCNN, gesd, margin_ranking, pad_sentences
        from mpit_b200.parallel.center import partition
        torch.backends.cuda.matmul.allow_tf32 = True
        self.gesd, self.margin_ranking = gesd, margin_ranking
        self.dev, self.rank, self.world, self.opt = dev, rank, world, opt
        torch.manual_seed(0)
        self.model = BiCNN(len(corpus.idx2word), 100, 200, 3000, 2, corpus.vectors).to(dev)
        self.ps = self.model.param_list()
        self.n = sum(p.numel() for p in self.ps)
        self.parts = partition(self.n, world)
        self.cap = max(s for _, s in self.parts)
        self.flat = torch.zeros(self.cap * world, device=dev)        # shard-major padded layout for the collectives
        self.gsend = torch.zeros(self.cap * world, device=dev)
        self.grecv = torch.zeros(self.cap * world, device=dev)
        self.m = torch.zeros(self.cap, device=dev)
        self.v = torch.zeros(self.cap, device=dev)
        self.t = 0
        qs, as_ = [s[1] for s in corpus.train], [s[2] for s in corpus.train]
        self.q_tok, self.q_len = pad_sentences(qs, dev)
        self.a_tok, self.a_len = pad_sentences(as_, dev)
        self.ans_tok, self.ans_len = pad_sentences(corpus.answers, dev)
        L = max(self.q_tok.shape[1], self.a_tok.shape[1], self.ans_tok.shape[1])
        pad = lambda t: torch.nn.functional.pad(t, (0, L - t.shape[1]))
        self.q_tok, self.a_tok, self.ans_tok = pad(self.q_tok), pad(self.a_tok), pad(self.ans_tok)
        self.targets = torch.tensor([s[0][0] for s in corpus.train], device=dev)
        self.A = len(corpus.answers)
        self.loss = torch.zeros((), device=dev)
        # position of flat element i in the shard-major padded layout of the collectives
        self.pos = torch.cat([torch.arange(sz, device=dev) + s * self.cap for s, (o, sz) in enumerate(self.parts)])

    def _flat_params(self):
        return torch.cat([p.detach().reshape(-1) for p in self.ps])

    def step(self, idx: torch.Tensor):
        opt, m, dev = self.opt, self.model, self.dev
        B, K = idx.numel(), opt["maxnegsample"]
        for p in self.ps:
            p.grad = None
        neg = torch.randint(0, self.A - 1, (B, K), device=dev)
        neg = neg + (neg >= self.targets[idx][:, None]).long()              # uniform over the non-target answers
        with torch.no_grad():
            q = m.encode(self.q_tok[idx], self.q_len[idx])
            sp = self.gesd(q, m.encode(self.a_tok[idx], self.a_len[idx]))
            cand = m.encode(self.ans_tok[neg.view(-1)], self.ans_len[neg.view(-1)])
            sn = self.gesd(q.repeat_interleave(K, 0), cand).view(B, K)
            viol = (sp[:, None] - sn) < opt["margin"]
            has = viol.any(1)
            first = viol.float().argmax(1)
        sel = neg.gather(1, first[:, None]).view(-1)
        qv = m.encode(self.q_tok[idx], self.q_len[idx])
        l = self.margin_ranking(self.gesd(qv, m.encode(self.a_tok[idx], self.a_len[idx])),
                                self.gesd(qv, m.encode(self.ans_tok[sel], self.ans_len[sel])), opt["margin"])
        loss = (l * has.float()).sum()
        loss.backward()
        self.loss = loss.detach()
        g = torch.cat([p.grad.reshape(-1) for p in self.ps])
        x = self._flat_params()
        g = (g + x * opt["L2reg"]).clamp_(-opt["gradClip"], opt["gradClip"])
        # push: every owner receives its shard of every client's gradient
        self.gsend[self.pos] = g
        self.flat[self.pos] = x
        if self.world > 1:
            dist.all_to_all_single(self.grecv, self.gsend)
        else:
            self.grecv.copy_(self.gsend)
        # server-side Adam, one client at a time (pserver.lua:140-155)
        mine = self.flat[self.rank * self.cap:(self.rank + 1) * self.cap].clone()
        b1, b2, lr, eps, sd = opt["beta1Adam"], opt["beta2Adam"], opt["lrAdam"], opt["epsilonAdam"], opt["stepDivAdam"]
        for c in range(self.world):
            gc = self.grecv[c * self.cap:(c + 1) * self.cap]
            self.t += 1
            k = self.t // sd + 1 if sd else self.t
            self.m.mul_(b1).add_(gc, alpha=1 - b1)
            self.v.mul_(b2).addcmul_(gc, gc, value=1 - b2)
            lr_t = lr * (1 - b2 ** k) ** 0.5 / (1 - b1 ** k)
            mine.addcdiv_(self.m, self.v.sqrt().add_(eps), value=-lr_t)
        # pull
        if self.world > 1:
            dist.all_gather_into_tensor(self.flat, mine)
        else:
            self.flat.copy_(mine)
        with torch.no_grad():
            newx = self.flat[self.pos]
            o = 0
            for p in self.ps:
                k = p.numel()
                p.copy_(newx[o:o + k].view_as(p))
                o += k


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("WORLD_SIZE", "1")))
    ap.add_argument("--impl", default="mpit", choices=["mpit", "torch"])
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--negs", type=int, default=100)
    ap.add_argument("--optimization", default="adam")
    ap.add_argument("--server", default="thread", choices=["thread", "inline"])
    ap.add_argument("--vocab", type=int, default=22354)
    ap.add_argument("--answers", type=int, default=2000)
    ap.add_argument("--train", type=int, default=2048)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from mpit_b200.launch.plaunch import parse
    from mpit_b200.workloads import bicnn as W
    from mpit_b200._native import launches
    opt = parse(["-optimization", a.optimization, "-batchSize", str(a.batch), "-maxnegsample", str(a.negs), "-type", "cuda"])
    corpus = make_corpus(W, a.vocab, a.answers, a.train)
    B = a.batch
    gen = torch.Generator().manual_seed(100 + rank)
    batches = [torch.randint(0, a.train, (B,), generator=gen).tolist() for _ in range(a.steps + a.warmup)]

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def max_over_ranks(v):
        t = torch.tensor([v], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    if a.impl == "mpit":
        from mpit_b200.parallel.symm import SymmetricHeap
        heap = SymmetricHeap(dev)
        tr = W.BiCNNDeviceTrainer(corpus, opt, dev, rank, heap, server=a.server)
        step = tr.train_step
        read_loss = lambda: float(tr.eng.loss_acc.item())
        nparams = tr.parameters.numel()
    else:
        tr = TorchPS(corpus, opt, dev, rank, world)

        def step(b):
            # a pinned block of its own per step (torch's pinned allocator recycles a block only after the copy that
            # reads it has run): the host may run ahead of the GPU
            tr.step(torch.tensor(b).pin_memory().to(dev, non_blocking=True))
        read_loss = lambda: float(tr.loss.item())
        nparams = tr.n
    barrier()
    for b in batches[:a.warmup]:
        step(b)
    barrier()
    # ---- device-timed arm: K steps between two events, no host read inside
    launches.count = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for b in batches[a.warmup:]:
        step(b)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = max_over_ranks(e0.elapsed_time(e1))
    n_launch = launches.count
    barrier()
    # ---- end-to-end arm: wall clock, indices from pinned memory every step, the loss read back every step
    t0 = time.perf_counter()
    for b in batches[a.warmup:]:
        step(b)
        read_loss()
    torch.cuda.synchronize(dev)
    wall = max_over_ranks(time.perf_counter() - t0)
    barrier()
    extra = {}
    if a.impl == "mpit":
        st = tr.loss_stats()
        extra = {"server": a.server, "drains": tr.close(), "applied_on_my_shard": tr.optim.ps.applied,
                 "max_flow_control_wait_us": round(tr.optim.ps.max_wait_us, 1), "triples": st["triples"],
                 "mean_hinge": round(st["hinge"] / max(st["steps"], 1), 5)}
        tr.optim.ps.check_timeout()
    if rank == 0:
        print(json.dumps({
            "metric": "questions/sec (device-timed, max over ranks) BiCNN parallel " + a.optimization, "impl": a.impl,
            "value": round(world * B * a.steps / (ms * 1e-3), 1), "unit": "questions/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": round(ms / a.steps, 4), "dtype": "tf32 (fp32 storage, fp32 accumulate)",
            "data": "synthetic corpus of the reference's shape, random-init weights",
            "config": {"model": f"BiCNN 100/200/3000 kw=2, V={a.vocab}, {nparams} params", "batch_per_gpu": B,
                       "candidates_per_question": a.negs, "sentences_per_step": B * (2 + a.negs),
                       "parallelism": f"async parameter server over NVLink, {world} shard owner(s)"},
            "e2e": {"value": round(world * B * a.steps / wall, 1), "unit": "questions/s",
                    "ms_per_step": round(wall * 1e3 / a.steps, 4), "h2d_bytes_per_step": 4 * B, "d2h_bytes_per_step": 4},
            "gpu_launches": n_launch, **extra}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
