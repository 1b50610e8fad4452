#!/usr/bin/env python
"""all-gather -> GEMM as ONE kernel (weights fetched from their owners' shards by the GEMM's own TMA loads) against
the un-fused formulations, device-timed, max over ranks.

    torchrun --nproc-per-node N tools/bench_pulled_linear.py [out.jsonl]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F

from mpit_b200._native import check, cur_stream, kernels, ptr
from mpit_b200.ops.pulled_linear import PulledLinear
from mpit_b200.parallel import SymmetricHeap

torch.backends.cuda.matmul.allow_tf32 = True
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
heap = SymmetricHeap(dev)
out_path = sys.argv[1] if len(sys.argv) > 1 else None
rows = []


def timed(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


for (M, K, N) in [(128, 1024, 4096), (128, 4096, 4096), (512, 1024, 4096), (1024, 1024, 4096), (128, 256, 2048)]:
    if N % (64 * world):
        continue
    lin = PulledLinear(heap, K, N, M, act=1)
    rpo = N // world
    lin.load_rows(torch.randn(rpo, K, device=dev) * 0.02)
    bias = torch.randn(N, device=dev) * 0.1
    x = torch.randn(M, K, device=dev)
    dy = torch.randn(M, N, device=dev) * 0.1
    lin.forward(x, bias)
    torch.cuda.synchronize()
    dist.barrier()

    _, ta_x = lin._operand(x, K, lin.x)

    def fused_one_tile_per_cta():            # the round-1 kernel: every M-tile fetches its weight tiles again
        check(kernels().mpit_umma_linear_pull(ta_x, lin._maps, lin.S, lin.rows_per_owner, lin._geom, ptr(bias), ptr(lin.y),
                                              ptr(lin.err), lin.block_n, cur_stream()), "mpit_umma_linear_pull")

    wfull = torch.empty(N, K, device=dev)
    wmine = lin.weight_rows()

    def nccl_then_cublas():
        dist.all_gather_into_tensor(wfull, wmine)
        return F.relu(F.linear(x, wfull, bias))

    def nccl_then_cublas_bwd():
        dist.all_gather_into_tensor(wfull, wmine)
        return dy @ wfull

    timed(lambda: lin.forward(x, bias), iters=10)          # clocks up before the first number that is kept
    t_f = timed(lambda: lin.forward(x, bias))
    t_1, t_u = timed(fused_one_tile_per_cta), timed(nccl_then_cublas)
    t_own = timed(lambda: lin._forward_unfused(ta_x, bias))
    t_g = timed(lambda: F.relu(F.linear(x, wfull, bias)))
    lin.forward(x, bias)
    ref = F.relu(F.linear(x, wfull, bias))
    rel = float((lin.y.view(lin.M, N)[:M] - ref).norm() / ref.norm())
    t_b, t_bu = timed(lambda: lin.backward(dy)), timed(nccl_then_cublas_bwd)
    # tiling sweeps (what the defaults were picked from): forward by M-tiles per CTA, backward by (M-tiles, split-K)
    m_tiles = (M + 127) // 128
    by_mt = {mt: round(timed(lambda mt=mt: lin.forward(x, bias, mt=mt), iters=20, warm=3), 1)
             for mt in (1, 2, 4) if mt == 1 or m_tiles >= mt}
    by_split = {f"{mt}x{ks}": round(timed(lambda mt=mt, ks=ks: lin.backward(dy, mt=mt, ksplit=ks), iters=20, warm=3), 1)
                for mt in (1, 4) if mt == 1 or m_tiles >= mt for ks in (1, 4, 8)}
    refb = dy @ wfull
    relb = float((lin.backward(dy) - refb).norm() / refb.norm())
    assert int(lin.err.item()) == 0
    r = {"M": M, "K": K, "N": N, "n_gpus": world, "weight_MB": round(N * K * 4 / 2 ** 20, 1), "fused_us": round(t_f, 1),
         "fused_one_tile_per_cta_us": round(t_1, 1), "own_gather_then_gemm_us": round(t_own, 1),
         "nccl_allgather_plus_cublas_us": round(t_u, 1), "cublas_gemm_alone_us": round(t_g, 1), "rel_err_vs_cublas": rel,
         "bwd_fused_us": round(t_b, 1), "bwd_nccl_allgather_plus_cublas_us": round(t_bu, 1), "bwd_rel_err_vs_cublas": relb,
         "fwd_us_by_mtiles_per_cta": by_mt, "bwd_us_by_mtiles_x_splitk": by_split}
    rows.append(r)
    if rank == 0:
        print(json.dumps(r), flush=True)
    del lin
heap.check_timeout()
if rank == 0 and out_path:
    with open(out_path, "w") as f:
        for r in rows:
            f.write(json.dumps(r) + "\n")
dist.barrier()
dist.destroy_process_group()
