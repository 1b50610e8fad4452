#!/usr/bin/env python
"""Headline benchmark: CIFAR-10-shaped 7-layer CNN, asynchronous EAMSGD (tau=10), one worker per
B200 with the centre variable sharded over the same GPUs (BASELINE.json config 3; reference
launcher asyncsgd/mlaunch.lua + asyncsgd/goot.lua).

    python bench.py --gpus 1 --steps 200 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0.  ``value`` is whole-job samples/s, device-timed (CUDA events) over
exactly K optimizer steps, max over ranks.  ``e2e`` is the same metric through the public
``CifarTrainer.train_step`` API including, every step, the host->device copy of that step's inputs
from pinned memory and a device->host read of the loss (wall clock, max over ranks).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

METRIC = "samples/sec (device-timed, max over ranks) CIFAR-10 CNN EAMSGD tau=10"
# multiply-accumulates of one forward pass per sample (asyncsgd/models/7-layers-cnn.lua): conv1 576*64*75,
# conv2 64*128*1600, conv3 4*64*1152, fc1 256*256, fc2 256*10
CONV1_MAC_PER_SAMPLE = 576 * 64 * 75
FLOP_FWD_MAC_PER_SAMPLE = CONV1_MAC_PER_SAMPLE + 64 * 128 * 1600 + 4 * 64 * 1152 + 256 * 256 + 256 * 10


def _clock_sampler(path):
    q = ("index,timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    try:
        f = open(path, "w")
        return subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "20"],
                                stdout=f, stderr=subprocess.DEVNULL), f
    except Exception:
        return None, None


def _clock_summary(path, dev_index, t_lo=None, t_hi=None):
    """Median SM clock and throttle reasons of the samples taken INSIDE the timed region [t_lo, t_hi] (datetime
    objects; nvidia-smi stamps every sample).  Falls back to all samples if the region was too short to catch one."""
    import datetime
    rows = []
    try:
        for line in open(path):
            p = [s.strip() for s in line.split(",")]
            if len(p) < 10 or not p[0].isdigit() or int(p[0]) != dev_index:
                continue
            try:
                ts = datetime.datetime.strptime(p[1], "%Y/%m/%d %H:%M:%S.%f")
            except ValueError:
                ts = None
            rows.append((ts, float(p[2]), float(p[3]), p[6:10]))
    except Exception:
        pass
    inside = [r for r in rows if r[0] is not None and t_lo is not None and t_lo <= r[0] <= t_hi]
    use = inside or rows
    sm, mx, reasons = [r[1] for r in use], max([r[2] for r in use], default=0.0), set()
    for r in use:
        for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3]):
            if v.lower().startswith("active"):
                reasons.add(name)
    return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx or None,
            "reasons": sorted(reasons), "samples": len(sm), "samples_inside_timed_region": len(inside)}


def run_reference_arm(args):
    """The unmodified reference is Lua/Torch7 + MPI; neither an interpreter nor MPI exists in this
    image and the tree has no Python packaging metadata (see DESIGN.md 'Reference arm')."""
    why = ("reference is Lua/Torch7 over MPI: pip install of /root/reference fails (no setup.py/pyproject), "
           "and no lua/luajit/th/mpirun/libmpi exists in this image")
    print(json.dumps({"impl": "reference", "unavailable": why, "n_gpus": args.gpus}))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=30)
    ap.add_argument("--impl", default="mpit", help="mpit (this framework) | reference | torch (library-call baseline)")
    ap.add_argument("--kernels", default=os.environ.get("MPIT_KERNELS", "auto"), help="umma | umma3x | ref | auto")
    ap.add_argument("--optim", default="eamsgd", help="eamsgd | easgd | downpour | sync_easgd | sync_eamsgd | sgd")
    ap.add_argument("--model", default="cifar", choices=["cifar", "mnist"],
                    help="cifar: the headline 7-layer CNN; mnist: the MNIST-shaped CNN of BASELINE.json config 2")
    ap.add_argument("--tau", type=int, default=10)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--no-graphs", action="store_true")
    ap.add_argument("--blocks", type=int, default=5, help="timed blocks of `steps` steps; the median block is reported")
    ap.add_argument("--no-fp32-arm", action="store_true", help="skip the 3xTF32 (fp32-accurate) arm")
    ap.add_argument("--no-baseline", action="store_true", help="skip the in-process library-call arm (vs_baseline = null)")
    ap.add_argument("--inline-loader", action="store_true",
                    help="e2e arm: assemble batches on the Python thread instead of the native loader thread")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.warmup < 3:
        args.warmup = 3
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    if args.impl == "torch":
        from baseline.torch_baseline import run as run_torch
        return run_torch(args, rank, world, dev)

    from mpit_b200._native import launches
    from mpit_b200.parallel.symm import SymmetricHeap
    from mpit_b200.workloads.cifar import CifarTrainer, CifarOpts, synthetic_cifar
    from mpit_b200.models.cifar_cnn import random_crop_offsets

    kernels = args.kernels
    if kernels == "auto":
        try:
            from mpit_b200.ops import umma_ops  # noqa: F401
            kernels = "umma"
        except Exception:
            kernels = "ref"
    B = args.batch
    opts = CifarOpts(name=args.optim, lr=1e-2, mom=0.99, mb=B, su=args.tau, mva=0.9 / world, impl=kernels,
                     use_graphs=not args.no_graphs, seed=1 + rank, model=args.model)
    CH = 1 if args.model == "mnist" else 3
    heap = SymmetricHeap(dev)
    tr = CifarTrainer(opts, dev, heap)

    # Synthetic CIFAR-shaped data, random-init weights.  Host set (pinned) for the e2e arm, and a
    # device-resident set LARGER than the 126 MB L2 for the kernel-only arm so every step's inputs
    # come from HBM.
    n_host = 64 * B
    hx, hy = synthetic_cifar(n_host, seed=100 + rank, channels=CH)
    n_dev_batches = 352 * 3 // CH            # 352 * 128 * 3072 B = 138 MB > L2
    dx = torch.randint(0, 256, (n_dev_batches, B, CH, 32, 32), dtype=torch.uint8, device=dev)
    dy = torch.randint(0, 10, (n_dev_batches, B), dtype=torch.int32, device=dev)
    dcrop = torch.randint(0, 5, (n_dev_batches, B, 2), dtype=torch.int32, device=dev)
    gen = torch.Generator().manual_seed(7 + rank)
    stage_x = torch.empty(B, CH, 32, 32, dtype=torch.uint8).pin_memory()
    stage_y = torch.empty(B, dtype=torch.int32).pin_memory()
    stage_c = torch.empty(B, 2, dtype=torch.int32).pin_memory()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def dev_step(i):
        j = i % n_dev_batches
        tr.engine.set_batch(dx[j], dy[j], dcrop[j])
        tr.device_step()

    # The e2e arm's batches come from the framework's loader: a native worker thread shuffles the host data set,
    # copies the samples of a mini-batch into a pinned slot and draws the crop offsets (goot.lua:164,172-184);
    # train_step then copies that slot host -> device.  --inline-loader assembles the batch on the Python thread
    # instead (the round-1 arrangement).
    loader = None
    if not args.inline_loader:
        from mpit_b200.data.loader import BatchLoader
        loader = BatchLoader(hx, hy, B, nslots=4, crop_range=5, seed=7 + rank)

    def e2e_step(i):
        if loader is not None:
            return tr.train_step(*loader.next())
        lo = (i * B) % n_host
        stage_x.copy_(hx[lo:lo + B])                 # what a loader does: assemble the batch on the host
        stage_y.copy_(hy[lo:lo + B])
        stage_c.copy_(random_crop_offsets(B, gen))
        return tr.train_step(stage_x, stage_y, stage_c)

    # ---- warm-up (covers every graph variant: sync / post-sync / plain steps) ----
    for i in range(max(args.warmup, 3 * args.tau + 3)):
        dev_step(i)
    barrier()

    # ---- kernel-only arm: device-timed.  BLOCKS back-to-back blocks of exactly `steps` steps, each bracketed by
    # barrier + synchronize; the reported block is the median one (a 20-step block lasts under 3 ms) ----
    BLOCKS = args.blocks
    clk_path = f"/tmp/mpit_clocks_{rank}.csv"
    sampler, fh = _clock_sampler(clk_path) if rank == 0 else (None, None)
    time.sleep(0.3 if sampler else 0.0)
    import datetime
    t_lo = datetime.datetime.now()
    dev_ms, launches_per_block = [], 0
    for b in range(BLOCKS):
        barrier()
        l0 = launches.count
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            dev_step(b * args.steps + i)
        e1.record()
        barrier()
        launches_per_block = launches.count - l0
        dev_ms.append(e0.elapsed_time(e1))
    loss_dev = float(tr.engine.loss_sum.item()) / B

    # ---- e2e arm: wall clock around the public API, H2D inputs + D2H loss every step ----
    for i in range(5):
        e2e_step(i)
    e2e_ms = []
    for b in range(BLOCKS):
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            e2e_step(i)
        last = tr.last_loss()
        barrier()
        e2e_ms.append((time.perf_counter() - t0) * 1e3)
    t_hi = datetime.datetime.now()
    if sampler:
        sampler.terminate()
        fh.close()
    heap.check_timeout()

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)          # per block: the slowest rank
    ms_max, e2e_ms_max = statistics.median(t[0].tolist()), statistics.median(t[1].tolist())
    gb = B * world
    value = gb * args.steps / (ms_max / 1e3)
    e2e_value = gb * args.steps / (e2e_ms_max / 1e3)

    # ---- the library-call formulation of the same step (cuDNN / cuBLAS / NCCL through torch), same process, same
    # box, same steps / warm-up / blocks and the same end-to-end rules: the anchor of `vs_baseline` ----
    base = None
    if not args.no_baseline and args.model == "cifar" and args.optim in ("eamsgd", "easgd"):
        try:
            del dx, dy, dcrop
            torch.cuda.empty_cache()
            from baseline.torch_baseline import measure as measure_torch
            base = measure_torch(args, rank, world, dev, BLOCKS)
        except Exception as ex:      # the product number must not depend on the baseline arm
            base = {"error": repr(ex)[:200]}

    # ---- the fp32-accurate arm: the same step with 3xTF32 products (the reference computes in fp32), device-timed the
    # same way, next to the libraries' fp32 kernels (allow_tf32 = False) ----
    fp32 = None
    if kernels == "umma" and args.model == "cifar" and not args.no_fp32_arm:
        try:
            opts3 = CifarOpts(name=args.optim, lr=1e-2, mom=0.99, mb=B, su=args.tau, mva=0.9 / world, impl="umma3x",
                              use_graphs=not args.no_graphs, seed=1 + rank)
            tr3 = CifarTrainer(opts3, dev, heap)
            dx3 = torch.randint(0, 256, (352, B, 3, 32, 32), dtype=torch.uint8, device=dev)
            dy3 = torch.randint(0, 10, (352, B), dtype=torch.int32, device=dev)
            dc3 = torch.randint(0, 5, (352, B, 2), dtype=torch.int32, device=dev)

            def step3(i):
                j = i % 352
                tr3.engine.set_batch(dx3[j], dy3[j], dc3[j])
                tr3.device_step()

            for i in range(max(args.warmup, 3 * args.tau + 3)):
                step3(i)
            ms3 = []
            for b in range(BLOCKS):
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(args.steps):
                    step3(b * args.steps + i)
                e1.record()
                barrier()
                ms3.append(e0.elapsed_time(e1))
            t3 = torch.tensor(ms3, dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t3, op=dist.ReduceOp.MAX)
            m3 = statistics.median(t3.tolist())
            fp32 = {"kernels": "umma3x", "dtype": "fp32-accurate: three tf32 tcgen05.mma per product on operands split "
                    "in shared memory, fp32 accumulate", "value": round(B * world * args.steps / (m3 / 1e3), 1),
                    "ms_per_step": round(m3 / args.steps, 5), "loss_last": round(float(tr3.engine.loss_sum.item()) / B, 4)}
            del tr3, dx3, dy3, dc3
            torch.cuda.empty_cache()
            if base and "value" in base:
                from baseline.torch_baseline import measure as measure_torch
                b32 = measure_torch(args, rank, world, dev, BLOCKS, tf32=False, e2e=False)
                fp32["library_fp32_value"] = round(b32["value"], 1)
                fp32["library_fp32_ms_per_step"] = round(b32["ms_per_step"], 5)
                fp32["ratio_vs_library_fp32"] = round(fp32["value"] / b32["value"], 3)
        except Exception as ex:
            fp32 = {"error": repr(ex)[:200]}

    if rank == 0:
        n_params = tr.engine.layout.n_flat
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")))
        except Exception:
            pass
        bf16 = float(peaks.get("bf16_tflops", 1641.8))
        flop_step = 2.0 * FLOP_FWD_MAC_PER_SAMPLE * B * 3 - 2.0 * CONV1_MAC_PER_SAMPLE * B   # fwd + dgrad + wgrad, no dgrad of conv1
        tflops = flop_step / (ms_max / args.steps * 1e-3) / 1e12
        vsb, detail = None, None
        if base and "value" in base:
            vsb = round(value / base["value"], 3)
            detail = {"impl": "torch: cuDNN/cuBLAS TF32 + NCCL all_gather/reduce_scatter, CUDA graph, same process",
                      "value": round(base["value"], 1), "e2e": round(base["e2e"], 1), "ratio": vsb,
                      "e2e_ratio": round(e2e_value / base["e2e"], 3), "ms_per_step": round(base["ms_per_step"], 5),
                      "e2e_ms_per_step": round(base["e2e_ms_per_step"], 5)}
        elif base:
            detail = base
        out = {
            "metric": METRIC if (args.model, args.optim, args.tau) == ("cifar", "eamsgd", 10) else
            f"samples/sec (device-timed, max over ranks) {args.model} CNN {args.optim} tau={args.tau}",
            "value": round(value, 1), "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_max / args.steps, 5),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": vsb, "vs_baseline_detail": detail,
            "dtype": {"umma": "tf32 (fp32 storage, fp32 accumulate)",
                      "umma3x": "fp32-accurate 3xTF32 (fp32 storage, three tf32 MMAs per product, fp32 accumulate)"}.get(kernels, "fp32"),
            "data": "synthetic CIFAR-10-shaped uint8 images, random-init weights",
            "impl": "mpit",
            "config": {"model": "CIFAR-10 7-layer CNN (asyncsgd/models/7-layers-cnn.lua, 351946 params)"
                       if args.model == "cifar" else f"MNIST-shaped CNN 1x32x32, 32/64/64 channels ({n_params} params)",
                       "global_batch": gb, "per_gpu_batch": B, "seq_len": None, "optimizer": args.optim,
                       "tau": args.tau, "lr": 1e-2, "mom": 0.99, "mva": 0.9 / world,
                       "parallelism": f"async-dp{world} + centre sharded over {world} GPU(s)",
                       "kernels": kernels, "cuda_graphs": not args.no_graphs,
                       "timing": f"median of {BLOCKS} back-to-back blocks of {args.steps} steps, each between barrier+synchronize",
                       "l2": "kernel-only arm rotates inputs through a 138 MB device-resident set (> 126 MB L2); no explicit flush"},
            "blocks_ms": [round(v, 4) for v in t[0].tolist()], "e2e_blocks_ms": [round(v, 4) for v in t[1].tolist()],
            "e2e": {"value": round(e2e_value, 1), "unit": "samples/s", "ms_per_step": round(e2e_ms_max / args.steps, 5),
                    "h2d_bytes_per_step": int(tr.h2d_bytes), "d2h_bytes_per_step": int(tr.d2h_bytes),
                    "timing": "wall clock around BatchLoader.next + CifarTrainer.train_step, max over ranks",
                    "loader": "inline (Python thread)" if loader is None else "native worker thread, pinned ring",
                    "host_path": "native stepper" if getattr(tr, "_stepper", None) is not None else "torch streams/events",
                    "d2h": "loss of step t stored into host-mapped pinned memory by the input kernel of step t+1 "
                           "(a copy for the newest step); read by the host every step, LAG=2 steps late"
                    if getattr(tr, "_stepper", None) is not None else "cudaMemcpyAsync of the loss every step"},
            "fp32_accurate": fp32,
            "gpu_launches": int(launches_per_block),
            "clocks": _clock_summary(clk_path, local_rank, t_lo, t_hi),
            "loss": {"device_arm_last": round(loss_dev, 4), "e2e_arm_last": round(last, 4)},
            "exchange_bytes_per_sync": 2 * 4 * n_params,
            "roofline": {"flop_per_step": flop_step, "achieved_tflops": round(tflops, 2),
                         "tf32_peak_tflops": round(bf16 / 2, 1), "tf32_peak_source": "half of MEASURED_PEAKS.json bf16_tflops",
                         "fraction_of_tf32_peak": round(tflops / (bf16 / 2), 4),
                         "hbm_gbs_measured": peaks.get("hbm_gbs"), "nvlink_gbs_per_direction": 900.0,
                         "exchange_us_at_nvlink_peak": round(2 * 4 * n_params * (world - 1) / max(world, 1) / 900e9 * 1e6 / 2, 3)},
        }
        if args.model != "cifar":
            out["roofline"] = None          # the FLOP count above is the CIFAR network's
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
