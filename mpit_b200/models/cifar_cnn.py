"""CIFAR-10 7-layer CNN (reference: asyncsgd/models/7-layers-cnn.lua:6-27, driven by
asyncsgd/goot.lua:154-322).

Architecture (identical to the reference, 351 946 parameters)::

    Conv(3->64,5x5) ReLU MaxPool2 | Conv(64->128,5x5) ReLU MaxPool2 | Conv(128->64,3x3) ReLU
    View(256) Linear(256->256) ReLU Dropout(0.5) | Linear(256->10) LogSoftMax + ClassNLL

What is different from the reference is *how* it runs.  The reference loops over the 128
samples of a mini-batch one at a time through 13 nn modules (goot.lua:198-210); here a whole
mini-batch is one explicit forward/backward schedule of fused kernels over NHWC activations,
captured into a CUDA graph, with:

* the flat parameter vector ``theta`` kept in the reference's ``getParameters()`` order
  (goot.lua:41) so that sharding, the elastic exchange and Torch7 checkpoints see exactly the
  reference layout, and
* a *packed* operand copy ``wpk`` (``[CO][KH][KWP][C]`` per layer, channel/kernel-width padded
  for 128-byte TMA rows) that the optimizer kernel refreshes in the same pass that updates
  ``theta``; gradients come back in the same packed layout and are gathered by the optimizer
  through ``wmap``.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional

import torch

from ..ops.nn_ops import ConvGeom

NUM_CLASSES = 10
IMG = 32      # stored image side (goot.lua:147-148)
CROP = 28     # training / test window (goot.lua:149-150)


@dataclass
class Layer:
    name: str
    cin: int
    cout: int
    k: int
    in_hw: int      # logical input side seen by the conv
    src_hw: int     # stored input side (differs from in_hw only for the cropped first layer)
    pool: bool
    relu: bool
    dropout: bool
    C: int          # physical input channels
    KWP: int        # physical kernel width (taps)
    CO: int         # physical output channels
    kw_taps: int = 0    # physical horizontal taps when they differ from k (pair-expanded first layer)
    # filled by build_layout
    w_off: int = 0
    b_off: int = 0
    d_off: int = -1     # data-gradient operand copy [C][KH][KW][CO] (-1: layer needs none)
    flat_w: int = 0
    flat_b: int = 0

    @property
    def out_hw(self):
        return self.in_hw - self.k + 1

    @property
    def act_hw(self):
        return self.out_hw // 2 if self.pool else self.out_hw

    def geom(self, B: int) -> ConvGeom:
        return ConvGeom(B, self.src_hw, self.src_hw, self.in_hw, self.in_hw, self.C, self.k,
                        self.kw_taps or self.k, self.KWP, self.out_hw, self.out_hw, self.CO)

    @property
    def wsize(self):
        return self.CO * self.k * self.KWP * self.C


def convnet_layers(cin: int = 3, c1: int = 64, c2: int = 128, c3: int = 64, f1: int = 256,
                   ncls: int = NUM_CLASSES) -> List[Layer]:
    """Layer table of the Conv5-pool / Conv5-pool / Conv3 / FC / FC topology of 7-layers-cnn.lua with
    free channel counts.  fc layers are 1x1 convolutions over a 1x1 'image'.  Constraints of the
    tensor-core tiles: cin <= 4 (pair-expanded 4-channel pixels), c1/c2/c3/f1 multiples of 32."""
    assert 1 <= cin <= 4 and all(c % 32 == 0 for c in (c1, c2, c3, f1)) and ncls <= 32
    return [
        # conv1 reads the pair-expanded image (8 floats per pixel = this pixel + the next one, 4
        # channels each): physically a 5x4-tap convolution over 8 channels whose 32-float rows hold
        # kernel columns [0,1 | 1,2 | 2,3 | 3,4]; the duplicated columns carry zero weights.
        Layer("conv1", cin, c1, 5, CROP, IMG, True, True, False, C=8, KWP=4, CO=c1, kw_taps=4),
        Layer("conv2", c1, c2, 5, 12, 12, True, True, False, C=c1, KWP=5, CO=c2),
        Layer("conv3", c2, c3, 3, 4, 4, False, True, False, C=c2, KWP=3, CO=c3),
        Layer("fc1", 4 * c3, f1, 1, 1, 1, False, True, True, C=4 * c3, KWP=1, CO=f1),
        Layer("fc2", f1, ncls, 1, 1, 1, False, False, False, C=f1, KWP=1, CO=32),
    ]


def cifar_layers() -> List[Layer]:
    """asyncsgd/models/7-layers-cnn.lua:8-24."""
    return convnet_layers(3, 64, 128, 64, 256, NUM_CLASSES)


def mnist_cnn_layers() -> List[Layer]:
    """MNIST-shaped CNN (1 input channel) for BASELINE.json config 2; see models/mnist.py."""
    return convnet_layers(1, 32, 64, 64, 256, NUM_CLASSES)


def _align(n: int, a: int = 32) -> int:
    return (n + a - 1) // a * a


class ParamLayout:
    """Flat (reference order) <-> packed (tensor-core operand order) index maps."""

    def __init__(self, layers: Optional[List[Layer]] = None):
        self.layers = layers or cifar_layers()
        flat = 0
        pk = 0
        for L in self.layers:
            L.flat_w = flat
            flat += L.cout * L.cin * L.k * L.k
            L.flat_b = flat
            flat += L.cout
            L.w_off = pk
            pk += _align(L.wsize)
            L.b_off = pk
            pk += _align(L.CO)
        for i, L in enumerate(self.layers):
            if i > 0:   # the first layer never needs a gradient w.r.t. its input
                L.d_off = pk
                pk += _align(L.C * L.k * L.k * L.CO)
        self.n_flat = flat
        self.n_packed = pk
        self.wmap = self._build_map()
        self.dmap = self._build_dmap()
        # stacked tables consumed by the optimizer kernels' operand refresh (pack_store)
        self.wmap2 = torch.cat([self.wmap, self.dmap])

    def _build_map(self) -> torch.Tensor:
        m = torch.empty(self.n_flat, dtype=torch.int64)
        for L in self.layers:
            co, ci, kh, kw = torch.meshgrid(torch.arange(L.cout), torch.arange(L.cin), torch.arange(L.k),
                                            torch.arange(L.k), indexing="ij")
            if L.name == "fc1":
                # Torch's View(64*2*2) flattens CHW; our activations are HWC
                # (7-layers-cnn.lua:19): flat input index i = c*4 + h*2 + w  ->  (h*2+w)*64 + c
                c, hw = ci // 4, ci % 4
                pc = hw * (L.cin // 4) + c
                tap = kw
            elif L.name == "conv1":
                # kernel column kw lives in slot [0,1,3,5,7][kw] of the 8 four-channel slots of a row
                slot = torch.tensor([0, 1, 3, 5, 7])[kw]
                tap = slot // 2
                pc = (slot % 2) * 4 + ci
            else:
                pc = ci
                tap = kw
            off = ((co * L.k + kh) * L.KWP + tap) * L.C + pc + L.w_off
            m[L.flat_w:L.flat_w + off.numel()] = off.reshape(-1)
            m[L.flat_b:L.flat_b + L.cout] = torch.arange(L.cout) + L.b_off
        assert m.unique().numel() == self.n_flat
        return m.to(torch.int32)

    def _build_dmap(self) -> torch.Tensor:
        """flat index -> slot in the data-gradient operand copy Wd[c][kh][kw][co] (K-major over
        (kh,kw,co)), -1 for parameters that have none (first layer, biases)."""
        m = torch.full((self.n_flat,), -1, dtype=torch.int64)
        for L in self.layers:
            if L.d_off < 0:
                continue
            co, ci, kh, kw = torch.meshgrid(torch.arange(L.cout), torch.arange(L.cin), torch.arange(L.k),
                                            torch.arange(L.k), indexing="ij")
            if L.name == "fc1":
                pc = (ci % 4) * (L.cin // 4) + ci // 4
            else:
                pc = ci
            off = ((pc * L.k + kh) * L.k + kw) * L.CO + co + L.d_off
            m[L.flat_w:L.flat_w + off.numel()] = off.reshape(-1)
        return m.to(torch.int32)

    def pack(self, theta: torch.Tensor) -> torch.Tensor:
        out = torch.zeros(self.n_packed, dtype=theta.dtype, device=theta.device)
        out[self.wmap.to(theta.device).long()] = theta
        dm = self.dmap.to(theta.device).long()
        sel = dm >= 0
        out[dm[sel]] = theta[sel]
        return out

    def unpack(self, packed: torch.Tensor) -> torch.Tensor:
        return packed[self.wmap.to(packed.device).long()]


# ---------------------------------------------------------------------------------------------
# Plain PyTorch restatement (CPU oracle, CPU workers, and the library-call GPU baseline)
# ---------------------------------------------------------------------------------------------
class TorchCifarCNN(torch.nn.Module):
    def __init__(self, cin: int = 3, c1: int = 64, c2: int = 128, c3: int = 64, f1: int = 256,
                 ncls: int = NUM_CLASSES):
        super().__init__()
        nn = torch.nn
        self.features = nn.Sequential(
            nn.Conv2d(cin, c1, 5), nn.ReLU(), nn.MaxPool2d(2, 2),
            nn.Conv2d(c1, c2, 5), nn.ReLU(), nn.MaxPool2d(2, 2),
            nn.Conv2d(c2, c3, 3), nn.ReLU())
        self.fc1 = nn.Linear(4 * c3, f1)
        self.drop = nn.Dropout(0.5)
        self.fc2 = nn.Linear(f1, ncls)
        torch7_reset(self)

    def forward(self, x, drop_mask=None):
        h = self.features(x).flatten(1)
        h = torch.relu(self.fc1(h))
        if drop_mask is not None:
            h = h * drop_mask
        else:
            h = self.drop(h)
        return torch.log_softmax(self.fc2(h), dim=1)

    def param_list(self):
        mods = [self.features[0], self.features[3], self.features[6], self.fc1, self.fc2]
        out = []
        for m in mods:
            out += [m.weight, m.bias]
        return out


def torch7_reset(model: torch.nn.Module, generator: Optional[torch.Generator] = None) -> None:
    """Torch7's default init: U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight and bias."""
    for m in model.modules():
        if isinstance(m, (torch.nn.Conv2d, torch.nn.Linear)):
            fan_in = m.weight[0].numel()
            bound = 1.0 / fan_in ** 0.5
            with torch.no_grad():
                m.weight.uniform_(-bound, bound, generator=generator)
                m.bias.uniform_(-bound, bound, generator=generator)


def flat_from_torch(model: TorchCifarCNN, grads: bool = False) -> torch.Tensor:
    ps = model.param_list()
    return torch.cat([(p.grad if grads else p.detach()).reshape(-1) for p in ps])


def flat_to_torch(model: TorchCifarCNN, theta: torch.Tensor) -> None:
    o = 0
    with torch.no_grad():
        for p in model.param_list():
            p.copy_(theta[o:o + p.numel()].view_as(p))
            o += p.numel()


# ---------------------------------------------------------------------------------------------
# Kernel engine
# ---------------------------------------------------------------------------------------------
class CifarEngine:
    """Batched forward/backward of the CIFAR CNN on one B200 through mpit_b200 kernels.

    ``impl='ref'`` uses the exact-fp32 CUDA-core kernels, ``impl='umma'`` the tcgen05/TMEM/TMA
    implicit-GEMM kernels (TF32 inputs, fp32 accumulate).  Both read ``wpk`` and write ``gpk``.
    """

    def __init__(self, batch: int, device="cuda", impl: str = "umma", seed: int = 1,
                 layers: Optional[List[Layer]] = None):
        from .. import ops
        self.ops = ops
        self.B = batch
        self.dev = torch.device(device)
        self.impl = impl
        self.layout = ParamLayout(layers)
        self.cin = self.layout.layers[0].cin
        self.ncls = self.layout.layers[-1].cout
        self.layers = self.layout.layers
        dev = self.dev
        f32 = dict(dtype=torch.float32, device=dev)
        B = batch
        self.wmap = self.layout.wmap2.to(dev)          # [2 * n_flat]: forward + data-gradient tables
        self.gmap = self.wmap[:self.layout.n_flat]     # gradients come back in the forward layout
        self.wmap_d = self.wmap[self.layout.n_flat:]   # flat index -> slot of the data-gradient operand layout
        # packed (forward layout) index -> flat index, for the optimizer epilogue of the weight-gradient reduction
        self.pk2flat = torch.full((self.layout.n_packed,), -1, dtype=torch.int32, device=dev)
        self.pk2flat[self.gmap.long()] = torch.arange(self.layout.n_flat, dtype=torch.int32, device=dev)
        self.opt_epi = None                            # set by an optimizer that wants the reduction to finish the step
        self.wpk = torch.zeros(self.layout.n_packed, **f32)
        self.gpk = torch.zeros(self.layout.n_packed, **f32)
        # inputs
        self.x_u8 = torch.zeros(B, self.cin, IMG, IMG, dtype=torch.uint8, device=dev)
        self.labels = torch.zeros(B, dtype=torch.int32, device=dev)
        self.crop = torch.zeros(B, 2, dtype=torch.int32, device=dev)
        self._labels_in = torch.zeros(B, dtype=torch.int32, device=dev)      # landing buffers for host inputs
        self._crop_in = torch.zeros(B, 2, dtype=torch.int32, device=dev)
        # pair-expanded NHWC8 image (+1 image of slack for the 4-pixel TMA rows at the image edge)
        self.x0 = torch.zeros((B + 1) * IMG * IMG * 8, **f32)
        # activations (NHWC)
        L1, L2, L3, L4, L5 = self.layers
        self.act = [torch.zeros(B * L.act_hw * L.act_hw * L.CO, **f32) for L in self.layers]
        self.idx = [torch.zeros(B * L.act_hw * L.act_hw * L.CO, dtype=torch.uint8, device=dev) if L.pool else None
                    for L in self.layers]
        # gradients wrt layer outputs (post-activation) and pre-activation (expanded)
        self.dact = [torch.zeros_like(a) for a in self.act]
        self.dz = [torch.zeros(B * L.out_hw * L.out_hw * L.CO, **f32) for L in self.layers]
        self.loss_sum = torch.zeros(1, **f32)
        self.confusion = torch.zeros(self.ncls, self.ncls, dtype=torch.int32, device=dev)
        self.rng = torch.tensor([seed, 0], dtype=torch.int64, device=dev)
        self._umma = None
        # umma3x: the same tcgen05 kernels issuing three MMAs per k-step on operands split in shared memory
        # (products of fp32 accuracy, the reference's precision; see mpit_set_tf32x3 in include/mpit_b200.h).
        # The mode is a per-launch property carried by a process-wide setting, so every schedule below sets it first
        # (captured graphs keep the mode they were captured with).
        from .._native import kernels as _kernels
        self.k = _kernels()
        self.x3 = {"umma": 0, "umma3x": int(os.environ.get("MPIT_X3_MODE", "1"))}.get(impl, 0)
        if impl in ("umma", "umma3x"):
            from ..ops import umma_ops
            self._umma = umma_ops.CifarUmmaPlan(self)

    # -- parameter plumbing --------------------------------------------------------------------
    def load_flat(self, theta: torch.Tensor) -> None:
        """Refresh the packed operand copy from a flat parameter vector."""
        self.ops.local_step(theta, None, None, None, None, self.wpk, self.wmap, do_apply=False)

    def flat_grad(self, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        out = out if out is not None else torch.empty(self.layout.n_flat, dtype=torch.float32, device=self.dev)
        self.ops.map_copy(out, self.gpk, self.gmap, scatter=False)
        return out

    def set_batch(self, images_u8: torch.Tensor, labels: torch.Tensor, crop: torch.Tensor) -> None:
        """Stage one mini-batch: ONE launch converts the uint8 images into the network's input layout
        (/255, pair-expanded NHWC8) and moves labels and crop offsets into the engine's buffers.  Device tensors
        are read in place (no staging copies: three small device-to-device copies in front of every step cost
        ~10 us of a 140 us step); host tensors are first copied (asynchronously when pinned -- the caller must then
        leave them alone until the stream has passed this point; ``CifarTrainer.train_step`` does that bookkeeping).  The conversion is
        NOT part of the captured step graph, so ``forward`` / ``fwd_bwd`` / ``evaluate`` always use the batch
        staged by the latest ``set_batch``."""
        if not images_u8.is_cuda:
            self.x_u8.copy_(images_u8, non_blocking=True)
            images_u8 = self.x_u8
        if not labels.is_cuda:
            self._labels_in.copy_(labels, non_blocking=True)
            labels = self._labels_in
        if not crop.is_cuda:
            self._crop_in.copy_(crop, non_blocking=True)
            crop = self._crop_in
        assert images_u8.dtype == torch.uint8 and images_u8.is_contiguous() and images_u8.numel() == self.x_u8.numel()
        assert labels.dtype == torch.int32 and crop.dtype == torch.int32 and labels.numel() == self.B \
            and crop.numel() == 2 * self.B and labels.is_contiguous() and crop.is_contiguous()
        self.ops.prep_u8(images_u8, self.x0, self.B, IMG, IMG, self.cin, labels=labels, labels_dst=self.labels,
                         crop=crop, crop_dst=self.crop)

    def w(self, i):
        L = self.layers[i]
        return self.wpk[L.w_off:L.w_off + L.wsize]

    def b(self, i):
        L = self.layers[i]
        return self.wpk[L.b_off:L.b_off + L.CO]

    def gw(self, i):
        L = self.layers[i]
        return self.gpk[L.w_off:L.w_off + L.wsize]

    def gb(self, i):
        L = self.layers[i]
        return self.gpk[L.b_off:L.b_off + L.CO]

    # -- schedules -----------------------------------------------------------------------------
    def forward(self, train: bool = True, classifier: bool = True) -> None:
        """All layers up to the logits, on the batch staged by :meth:`set_batch`.  ``classifier=False`` (used by
        :meth:`fwd_bwd` / :meth:`evaluate` when the fused classifier head will follow) stops after the last hidden
        layer."""
        ops, B = self.ops, self.B
        if self._umma is not None:
            self.k.mpit_set_tf32x3(self.x3)
            self._umma.forward(train, classifier)
        else:
            src = self.x0
            for i, L in enumerate(self.layers):
                ops.conv_fwd_ref(L.geom(B), src, self.w(i), self.b(i), self.act[i],
                                 crop=self.crop if i == 0 else None, pool_idx=self.idx[i], relu=L.relu,
                                 pool=L.pool, dropout=(L.dropout and train), rng=self.rng)
                src = self.act[i]

    def _fused_head(self, train: bool) -> bool:
        um = self._umma
        return um is not None and um.head and (not train or um.wgrad is not None)

    def loss_and_grad(self, train: bool = True) -> None:
        L5 = self.layers[-1]
        if self._fused_head(train) and self._umma.head_loss(train):
            return
        # with the tensor-core plan the classifier's bias gradient is produced here (the other layers' come from
        # the data-gradient epilogues); the reference path column-sums every dz in backward()
        fused_bias = train and self._umma is not None and self._umma.wgrad is not None
        self.ops.nll_fwd_bwd(self.act[-1], self.labels, self.dz[-1] if train else None, self.loss_sum,
                             self.B, self.ncls, L5.CO, gscale=1.0 / self.B, confusion=self.confusion,
                             rng_inc=self.rng if train else None, dbias=self.gb(len(self.layers) - 1) if fused_bias else None)

    def backward(self) -> None:
        ops, B = self.ops, self.B
        if self._umma is not None:
            self.k.mpit_set_tf32x3(self.x3)
            self._umma.backward()
            return
        for i in range(len(self.layers) - 1, -1, -1):
            L = self.layers[i]
            g = L.geom(B)
            x_in = self.act[i - 1] if i > 0 else self.x0
            dz = self.dz[i]
            ops.conv_wgrad_ref(g, dz, x_in, self.gw(i), crop=self.crop if i == 0 else None)
            ops.colsum(dz, self.gb(i), B * L.out_hw * L.out_hw, L.CO)
            if i == 0:
                break
            P = self.layers[i - 1]
            ops.conv_dgrad_ref(g, dz, self.w(i), self.dact[i - 1])
            ops.act_bwd(self.dact[i - 1], self.act[i - 1], self.dz[i - 1], B, P.out_hw, P.out_hw, P.CO,
                        pool_idx=self.idx[i - 1], pool=P.pool, scale=2.0 if P.dropout else 1.0)

    def fwd_bwd(self) -> None:
        """One training forward+backward: fills ``gpk`` and ``loss_sum`` (sum of per-sample NLL)."""
        self.forward(True, classifier=not self._fused_head(True))
        self.loss_and_grad(True)
        self.backward()

    def evaluate(self) -> None:
        """Forward only (no dropout); accumulates ``loss_sum`` and ``confusion`` (goot.lua:265-322)."""
        self.forward(False, classifier=not self._fused_head(False))
        self.loss_and_grad(False)

    def logits(self) -> torch.Tensor:
        return self.act[-1].view(self.B, self.layers[-1].CO)[:, :self.ncls]


def center_crop_offsets(B: int) -> torch.Tensor:
    """goot.lua:283-285: w1 = ceil((32-28)/2) = 2."""
    return torch.full((B, 2), (IMG - CROP + 1) // 2, dtype=torch.int32)


def random_crop_offsets(B: int, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """goot.lua:177: x1,y1 = random(0, 32-28) per sample; stored as (y, x)."""
    return torch.randint(0, IMG - CROP + 1, (B, 2), generator=generator, dtype=torch.int32)
