"""CIFAR-10 training/testing loop on B200s (reference: asyncsgd/goot.lua).

``CifarTrainer`` is the public entry point a user (and ``bench.py``) drives: it owns the kernel
engine, the device optimizer and -- for the distributed optimizers -- this rank's handle on the
NVLink-sharded centre variable.  ``train_step`` takes one mini-batch in pinned host memory, exactly
what a data loader hands over, and returns the (one step delayed) loss.

Differences from goot.lua that are deliberate: the mini-batch is one batched kernel schedule instead
of 128 sequential single-sample passes (goot.lua:198-210); random crops are applied by the first
convolution's TMA box coordinates instead of ``image.crop`` on the CPU (goot.lua:177-178); the
confusion matrix is accumulated on the device (goot.lua:209).
"""
from __future__ import annotations

import os
import time
from dataclasses import dataclass, field
from typing import Optional

import torch

from ..models.cifar_cnn import (CifarEngine, TorchCifarCNN, flat_from_torch, random_crop_offsets,
                                center_crop_offsets, NUM_CLASSES, IMG, CROP)
from ..optim.device import DeviceMSGD, DeviceEAMSGD, DeviceDownpour, DeviceRule, DeviceSyncEASGD
from ..parallel.center import ShardedCenter
from ..parallel.symm import SymmetricHeap


@dataclass
class CifarOpts:
    """Mirror of the ``opt`` table of goot.lua:6-21 plus launcher settings (mlaunch.lua:48-67)."""
    name: str = "eamsgd"          # sgd | downpour | eamsgd | easgd | sync_easgd | sync_eamsgd | <rule>single
    lr: float = 1e-2
    mom: float = 0.99
    mb: int = 128
    mva: float = 0.0              # moving rate alpha; default 0.9/p as in mlaunch.lua:57-58
    su: int = 1                   # communication period tau
    l2wd: float = 0.0
    lrd: float = 0.0
    lrp: float = 0.0
    maxepoch: int = 1
    saveep: int = 10
    seed: int = 1                 # per-worker: data order, crops, dropout stream
    init_seed: int = 1            # model initialisation: the same on every worker (goot.lua:29-30 seeds all ranks alike)
    impl: str = "umma"            # kernel path: umma (tcgen05) | ref (fp32 CUDA cores)
    use_graphs: bool = True
    save: str = "results"
    rule_hyper: dict = field(default_factory=dict)
    model: str = "cifar"          # cifar (3x32x32, 64/128/64 channels) | mnist (1x32x32, 32/64/64 channels)


def synthetic_cifar(n: int, seed: int = 0, pinned: bool = True, channels: int = 3):
    """Synthetic data of the CIFAR-10 shape: uint8 [n,3,32,32] images, labels in [0,10)
    (``channels=1`` gives the MNIST-shaped variant)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (n, channels, IMG, IMG), dtype=torch.uint8, generator=g)
    y = torch.randint(0, NUM_CLASSES, (n,), dtype=torch.int32, generator=g)
    if pinned and torch.cuda.is_available():
        x, y = x.pin_memory(), y.pin_memory()
    return x, y


def learnable_cifar(n: int, seed: int = 0, pinned: bool = True, channels: int = 3, noise: float = 40.0, proto_seed: int = 1234):
    """A synthetic set that CAN be learned: every class has its own smooth mean image (8x8 blocks of a random
    pattern fixed by ``proto_seed``) and a sample is that image plus pixel noise -- for convergence tests, where
    the uniform-noise set of :func:`synthetic_cifar` has nothing to fit."""
    gp = torch.Generator().manual_seed(proto_seed)
    proto = torch.randint(40, 216, (NUM_CLASSES, channels, 4, 4), generator=gp).float()
    proto = proto.repeat_interleave(IMG // 4, 2).repeat_interleave(IMG // 4, 3)           # [10, C, 32, 32]
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, NUM_CLASSES, (n,), dtype=torch.int32, generator=g)
    x = (proto[y.long()] + torch.randn(n, channels, IMG, IMG, generator=g) * noise).clamp_(0, 255).to(torch.uint8)
    if pinned and torch.cuda.is_available():
        x, y = x.pin_memory(), y.pin_memory()
    return x, y


def cifar_data(trsize: int, tesize: int, seed: int = 0, channels: int = 3, root: Optional[str] = None,
               pinned: bool = True):
    """(train_x, train_y), (test_x, test_y), is_real: the Torch7 CIFAR-10 batches the reference loads
    (goot.lua:62-77) when they can be found (``root`` / $CIFAR10_T7 / ./cifar-10-batches-t7), else synthetic
    data of the same shape.  Every worker sees the whole set and shuffles with its own seed, like the reference."""
    from ..data import find_cifar10, load_cifar10
    if channels == 3 and find_cifar10(root) is not None:
        (x, y), (tx, ty), _ = load_cifar10(root, trsize, tesize, synthetic_ok=False)
        if pinned and torch.cuda.is_available():
            x, y, tx, ty = x.pin_memory(), y.pin_memory(), tx.pin_memory(), ty.pin_memory()
        return (x, y), (tx, ty), True
    return (synthetic_cifar(trsize, seed=seed, pinned=pinned, channels=channels),
            synthetic_cifar(tesize, seed=2, pinned=pinned, channels=channels), False)


def load_checkpoint(path: str) -> torch.Tensor:
    """Flat parameters from a checkpoint written by :func:`train_epochs` (a Torch7 ``nn.Sequential``,
    goot.lua:246-254) or from a flat-tensor file as BiCNN writes (bicnn.lua:590-594) -- the resume path the
    reference only has for BiCNN (``-loadmodel``, bicnn.lua:259-261)."""
    from .. import t7
    obj = t7.load(path)
    if isinstance(obj, torch.Tensor):
        return obj.reshape(-1).float()
    return t7.cifar_theta_from_t7(obj)


class CifarTrainer:
    def __init__(self, opts: CifarOpts, device: Optional[torch.device] = None, heap: Optional[SymmetricHeap] = None,
                 theta0: Optional[torch.Tensor] = None):
        self.opts = opts
        self.dev = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        torch.cuda.set_device(self.dev)
        from ..models.cifar_cnn import mnist_cnn_layers, cifar_layers
        layers = mnist_cnn_layers() if opts.model == "mnist" else cifar_layers()
        self.engine = CifarEngine(opts.mb, device=self.dev, impl=opts.impl, seed=opts.seed, layers=layers)
        if theta0 is None:
            g = torch.Generator().manual_seed(opts.init_seed)
            m = TorchCifarCNN(1, 32, 64, 64, 256) if opts.model == "mnist" else TorchCifarCNN()
            from ..models.cifar_cnn import torch7_reset
            torch7_reset(m, g)
            theta0 = flat_from_torch(m)
        self.heap = heap
        self.center: Optional[ShardedCenter] = None
        name = opts.name
        n = theta0.numel()
        if name in ("sync_easgd", "sync_eamsgd") and heap is None:
            self.heap = heap = SymmetricHeap(self.dev)
        if name in ("eamsgd", "easgd", "downpour"):
            if heap is None:
                self.heap = heap = SymmetricHeap(self.dev)
            self.center = ShardedCenter.create(heap, n)
        p = heap.nranks if heap is not None else 1
        if name in ("eamsgd", "easgd"):
            mva = opts.mva if opts.mva > 0 else 0.9 / p
            self.opt = DeviceEAMSGD(self.engine, theta0, self.center, lr=opts.lr, mva=mva, su=opts.su,
                                    mom=0.0 if name == "easgd" else opts.mom, l2wd=opts.l2wd, lrd=opts.lrd,
                                    lrp=opts.lrp, use_graphs=opts.use_graphs)
        elif name in ("sync_easgd", "sync_eamsgd"):
            mva = opts.mva if opts.mva > 0 else 0.9
            self.opt = DeviceSyncEASGD(self.engine, theta0, heap, lr=opts.lr, mva=mva, su=opts.su,
                                       mom=0.0 if name == "sync_easgd" else opts.mom, l2wd=opts.l2wd,
                                       use_graphs=opts.use_graphs)
        elif name == "downpour":
            self.opt = DeviceDownpour(self.engine, theta0, self.center, lr=opts.lr, su=opts.su, l2wd=opts.l2wd,
                                      lrd=opts.lrd, use_graphs=opts.use_graphs)
        elif name in ("sgd", "msgd"):
            self.opt = DeviceMSGD(self.engine, theta0, lr=opts.lr, mom=opts.mom, l2wd=opts.l2wd, lrd=opts.lrd,
                                  lrp=opts.lrp, use_graphs=opts.use_graphs)
        elif name.endswith("single"):
            self.opt = DeviceRule(self.engine, theta0, name[:-6], opts.rule_hyper, self.center)
        else:
            raise ValueError(f"unknown optimizer {name}")
        if self.heap is not None and self.heap.nranks > 1:
            import torch.distributed as dist
            torch.cuda.synchronize()
            dist.barrier()   # centre seeded before anyone syncs against it
        # pinned landing zone for the per-step loss read-back
        # Pipelined host side of train_step: the loss comes back LAG steps late (so the host never waits for the
        # step it has just launched) and the inputs of step t travel on a copy stream into a double-buffered device
        # staging area while step t-1 computes.
        self.LAG = 2
        self._loss_host = torch.zeros(self.LAG + 1, 2, dtype=torch.float32).pin_memory()
        self._loss_dev = torch.zeros(self.LAG + 1, 2, dtype=torch.float32, device=self.dev)
        self._loss_evt = [torch.cuda.Event() for _ in range(self.LAG + 1)]
        self._k = 0                                   # steps issued through train_step
        self._ret = 0                                 # steps whose loss has been handed back
        e = self.engine
        self._copy_stream = torch.cuda.Stream(device=self.dev)
        self._stage = [(torch.empty_like(e.x_u8), torch.empty_like(e.labels), torch.empty_like(e.crop))
                       for _ in range(2)]
        self._in_ready = [torch.cuda.Event() for _ in range(2)]     # H2D of the slot has landed
        self._in_free = [torch.cuda.Event() for _ in range(2)]      # the slot has been consumed by the main stream
        self.tm = {"feval": 0.0, "sync": 0.0}
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        # The same choreography issued from one C function per half (csrc/kernels/stepper.cu) instead of a dozen
        # torch stream / event / copy calls: the Python calls cost about as much host time as the device step takes.
        # MPIT_NATIVE_STEPPER=0 keeps the torch path (same semantics).
        self._stepper = None
        if os.environ.get("MPIT_NATIVE_STEPPER", "1") != "0" and self.dev.type == "cuda":
            import ctypes
            from .._native import kernels, check, ptr
            from ..models.cifar_cnn import IMG
            h = ctypes.c_void_p(0)
            (x0, y0, c0), (x1, y1, c1) = self._stage
            with torch.cuda.device(self.dev):
                check(kernels().mpit_stepper_create(ctypes.byref(h), self.LAG, ptr(x0), ptr(x1), ptr(y0), ptr(y1),
                                                    ptr(c0), ptr(c1), ptr(self._loss_host), x0.numel(), opts.mb, IMG,
                                                    IMG, e.cin), "mpit_stepper_create")
            self._stepper = h
            self._prev = (ctypes.c_float(0.0), ctypes.c_int(0))

    def __del__(self):
        if getattr(self, "_stepper", None):
            try:
                from .._native import kernels
                kernels().mpit_stepper_destroy(self._stepper)
            except Exception:
                pass
            self._stepper = None

    def _train_step_native(self, images_u8, labels, crop) -> Optional[float]:
        import ctypes
        from .._native import kernels, check, ptr, cur_stream
        e, lib = self.engine, kernels()
        assert images_u8.dtype == torch.uint8 and images_u8.is_contiguous() and images_u8.numel() == e.x_u8.numel() \
            and labels.dtype == torch.int32 and crop.dtype == torch.int32 and labels.numel() == self.opts.mb \
            and crop.numel() == 2 * self.opts.mb and not images_u8.is_cuda
        prev, have = self._prev
        main = cur_stream()
        check(lib.mpit_stepper_stage(self._stepper, ptr(images_u8), ptr(labels), ptr(crop), ptr(e.x0), ptr(e.labels),
                                     ptr(e.crop), main, ctypes.byref(prev), ctypes.byref(have)), "mpit_stepper_stage")
        self.opt.step()
        rc = lib.mpit_stepper_finish(self._stepper, ptr(e.loss_sum), main)      # no kernel: not a counted launch
        if rc != 0:
            raise RuntimeError(f"mpit_stepper_finish failed with code {rc}")
        self._k += 1
        self.h2d_bytes = images_u8.numel() + labels.numel() * 4 + crop.numel() * 4
        self.d2h_bytes = 4
        if have.value:
            self._ret += 1
            return prev.value / self.opts.mb
        return None

    def _native_loss(self, index: int) -> float:
        import ctypes
        from .._native import kernels
        out = ctypes.c_float(0.0)
        rc = kernels().mpit_stepper_loss(self._stepper, index, ctypes.byref(out))
        if rc != 0:
            raise RuntimeError(f"mpit_stepper_loss({index}) failed with code {rc}")
        return out.value / self.opts.mb

    # -- one training step ---------------------------------------------------------------------
    def train_step(self, images_u8: torch.Tensor, labels: torch.Tensor, crop: torch.Tensor) -> Optional[float]:
        """Stage a mini-batch (pinned host tensors), run one optimizer step, and return the mean loss of the step
        issued ``LAG`` calls earlier (None for the first calls); :meth:`last_loss` gives the newest one.

        The caller may reuse its host buffers as soon as this returns: the host->device copies run on a copy
        stream (beside the previous step's kernels) and are waited for here; only the ~2 us device-to-device move
        into the engine's input buffers sits on the compute stream."""
        if self._stepper is not None:
            if not (images_u8.is_cuda or labels.is_cuda or crop.is_cuda):
                return self._train_step_native(images_u8, labels, crop)
            if self._k:
                raise ValueError("train_step: device-resident inputs after host inputs; use one kind per trainer")
            self.__del__()                       # device-resident inputs: the torch path copies device-to-device
        k, lag = self._k, self.LAG
        prev = None
        if k - self._ret >= lag:
            slot = self._ret % (lag + 1)
            self._loss_evt[slot].synchronize()
            prev = float(self._loss_host[slot, 0]) / self.opts.mb
            self._ret += 1
        main, cs, b = torch.cuda.current_stream(), self._copy_stream, k & 1
        sx, sy, sc = self._stage[b]
        if k >= 2:
            cs.wait_event(self._in_free[b])
        with torch.cuda.stream(cs):
            sx.copy_(images_u8, non_blocking=True)
            sy.copy_(labels, non_blocking=True)
            sc.copy_(crop, non_blocking=True)
            self._in_ready[b].record(cs)
        self.h2d_bytes = images_u8.numel() + labels.numel() * 4 + crop.numel() * 4
        main.wait_event(self._in_ready[b])
        self.engine.set_batch(sx, sy, sc)                  # device-to-device, on the compute stream
        self._in_free[b].record(main)
        self.opt.step()
        slot = k % (lag + 1)
        # straight from the engine's accumulator: the next step's loss kernel is stream-ordered after this copy
        self._loss_host[slot, :1].copy_(self.engine.loss_sum, non_blocking=True)
        self._loss_evt[slot].record()
        self.d2h_bytes = 4
        self._k = k + 1
        self._in_ready[b].synchronize()                    # the caller's host buffers are free again
        return prev

    def device_step(self) -> None:
        """One optimizer step on whatever batch is already staged on the device (kernel-only timing)."""
        self.opt.step()

    def drain_losses(self) -> list:
        """Mean losses of the steps :meth:`train_step` has not handed back yet (waits for them)."""
        out = []
        if self._stepper is not None:
            import ctypes
            from .._native import kernels
            out = [self._native_loss(i) for i in range(self._ret, self._k)]
            a, b = ctypes.c_longlong(0), ctypes.c_longlong(0)
            kernels().mpit_stepper_counters(self._stepper, ctypes.byref(a), ctypes.byref(b), 1)
            self._ret = self._k
            return out
        while self._ret < self._k:
            slot = self._ret % (self.LAG + 1)
            self._loss_evt[slot].synchronize()
            out.append(float(self._loss_host[slot, 0]) / self.opts.mb)
            self._ret += 1
        return out

    def last_loss(self) -> float:
        """Mean loss of the most recent step issued through :meth:`train_step` (waits for it)."""
        if self._stepper is not None:
            return self._native_loss(-1)
        slot = (self._k - 1) % (self.LAG + 1)
        self._loss_evt[slot].synchronize()
        return float(self._loss_host[slot, 0]) / self.opts.mb

    def params(self) -> torch.Tensor:
        return self.opt.params()

    # -- evaluation (goot.lua:265-322) ---------------------------------------------------------
    @torch.no_grad()
    def test(self, images_u8: torch.Tensor, labels: torch.Tensor) -> dict:
        eng, B = self.engine, self.opts.mb
        eng.load_flat(self.params())
        eng.confusion.zero_()
        total, n = 0.0, (images_u8.shape[0] // B) * B
        crop = center_crop_offsets(B)
        for i in range(0, n, B):
            eng.set_batch(images_u8[i:i + B], labels[i:i + B], crop)
            eng.evaluate()
            total += float(eng.loss_sum.item())
        conf = eng.confusion.cpu()
        return {"loss": total / max(n, 1), "acc": float(conf.diag().sum()) / max(int(conf.sum()), 1),
                "confusion": conf, "n": n}


def train_epochs(tr: CifarTrainer, train_xy, test_xy, rank: int = -1, quiet: bool = False) -> dict:
    """Epoch loop of goot.lua:154-261 around a :class:`CifarTrainer`: per-epoch shuffle, random crops,
    ``train.log``/``test.log`` rows, a Torch7 ``model-epoch-N.net`` every ``saveep`` epochs (a loadable
    ``nn.Sequential``, goot.lua:246-254), test after every epoch, timing summary (goot.lua:330-337)."""
    from .. import t7
    from ..utils import Logger
    opts = tr.opts
    B = opts.mb
    (trx, try_), (tex, tey) = train_xy, test_xy
    trsize = (trx.shape[0] // B) * B
    os.makedirs(opts.save, exist_ok=True)
    sfx = "" if rank < 0 else str(rank)
    train_log = Logger(os.path.join(opts.save, f"train{sfx}.log"))
    test_log = Logger(os.path.join(opts.save, f"test{sfx}.log"))
    gen = torch.Generator().manual_seed(opts.seed)
    stage_x = torch.empty(B, tr.engine.cin, IMG, IMG, dtype=torch.uint8).pin_memory()
    stage_y = torch.empty(B, dtype=torch.int32).pin_memory()
    stage_c = torch.empty(B, 2, dtype=torch.int32).pin_memory()
    hist = {"train_loss": [], "train_acc": [], "test_acc": [], "test_loss": []}
    t0 = time.perf_counter()
    n_steps = 0
    # Mini-batches come from the native loader thread (per-epoch permutation, crop offsets, copies into pinned
    # slots: goot.lua:164,172-184 off the training thread); MPIT_PY_LOADER=1 assembles them here instead.
    loader = None
    if os.environ.get("MPIT_PY_LOADER", "0") != "1" and not trx.is_cuda and trx.dtype == torch.uint8 \
            and try_.dtype == torch.int32 and trsize >= B:
        from ..data.loader import BatchLoader
        loader = BatchLoader(trx[:trsize].contiguous(), try_[:trsize].contiguous(), B, nslots=4,
                             crop_range=IMG - CROP + 1, seed=opts.seed)
    for epoch in range(1, opts.maxepoch + 1):
        shuffle = torch.randperm(trsize, generator=gen) if loader is None else None
        tr.engine.confusion.zero_()
        losses = []
        for t in range(0, trsize, B):
            if loader is not None:
                prev = tr.train_step(*loader.next())
            else:
                idx = shuffle[t:t + B]
                torch.index_select(trx, 0, idx, out=stage_x)
                torch.index_select(try_, 0, idx, out=stage_y)
                stage_c.copy_(random_crop_offsets(B, gen))
                prev = tr.train_step(stage_x, stage_y, stage_c)
            if prev is not None:
                losses.append(prev)
            n_steps += 1
        losses.extend(tr.drain_losses())
        conf = tr.engine.confusion.cpu()
        acc = float(conf.diag().sum()) / max(int(conf.sum()), 1)
        elapsed = (time.perf_counter() - t0) / 60
        train_log.add({"% train accuracy": f"{acc:0.4f}", "% wall-clock elapsed time(min)": f"{elapsed:0.2f}",
                       "% train loss": f"{sum(losses) / len(losses):0.4f}"})
        hist["train_loss"].append(sum(losses) / len(losses))
        hist["train_acc"].append(acc)
        if epoch % opts.saveep == 0:
            t7.save(os.path.join(opts.save, f"model-epoch-{epoch}.net"), t7.cifar_model_to_t7(tr.params()))
        res = tr.test(tex, tey)
        test_log.add({"% test accuracy": f"{res['acc']:0.4f}",
                      "% wall-clock elapsed time(min)": f"{(time.perf_counter() - t0) / 60:0.2f}",
                      "% test error": f"{res['loss']:0.2f}"})
        hist["test_acc"].append(res["acc"])
        hist["test_loss"].append(res["loss"])
    total = time.perf_counter() - t0
    hist["total"] = total
    hist["samples_per_s"] = n_steps * B / total
    if not quiet:
        print(rank, "total training time is", total)
        print(f"\n===> time to learn 1 sample = {total / max(n_steps * B, 1) * 1000:.5f}ms")
    train_log.close()
    test_log.close()
    if loader is not None:
        loader.close()
    return hist
