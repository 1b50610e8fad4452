"""tcgen05/TMEM/TMA layer kernels (csrc/kernels/umma_igemm.cu, umma_wgrad.cu): tensor-map
construction and the per-layer launch plans of the CIFAR CNN.

A plan is built once per engine: every TMA descriptor (activation boxes, weight boxes) is encoded
on the host up front and the launches carry them by value, so a training step is a fixed sequence
of kernel launches that captures cleanly into a CUDA graph.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Optional

import torch

from .._native import kernels, check, ptr, cur_stream


def make_tmap(t_ptr: int, dims: List[int], strides_bytes: List[int], box: List[int], swizzle128: int = 1):
    """Encode a tiled fp32 tensor map; dims/box innermost first, strides for dims 1..rank-1."""
    rank = len(dims)
    buf = (ctypes.c_ubyte * 128)()
    d = (ctypes.c_ulonglong * 5)(*dims)
    s = (ctypes.c_ulonglong * 5)(*strides_bytes)
    b = (ctypes.c_uint * 5)(*box)
    check(kernels().mpit_tmap_encode_f32(buf, ctypes.c_void_p(t_ptr), rank, d, s, b, int(swizzle128)),
          "mpit_tmap_encode_f32")
    return buf


@dataclass
class IgemmLaunch:
    tmap_a: object
    tmap_b: object
    geom: object          # c_int * 17
    scale: float
    bias: Optional[torch.Tensor]
    out: torch.Tensor
    idx: Optional[torch.Tensor]
    aux: Optional[torch.Tensor]
    aux_idx: Optional[torch.Tensor]
    crop: Optional[torch.Tensor]
    rng: Optional[torch.Tensor]
    block_n: int
    dropout_slot: int = 16
    colpart: Optional[torch.Tensor] = None    # [m_tiles][N] per-tile column sums (modes MASK / EXPAND)
    x3: int = 0                               # 0 TF32 products, 1 / 2 fp32-accurate 3xTF32 (mpit_set_tf32x3)

    def run(self, err: torch.Tensor, dropout: Optional[bool] = None):
        kernels().mpit_set_tf32x3(self.x3)    # the mode is process-wide: every launch states the one it wants
        g = self.geom
        if dropout is not None:
            g = (ctypes.c_int * 17)(*self.geom)
            g[16] = int(dropout)
        check(kernels().mpit_umma_igemm(self.tmap_a, self.tmap_b, g, self.scale, ptr(self.bias), ptr(self.out),
                                        ptr(self.idx), ptr(self.aux), ptr(self.aux_idx), ptr(self.crop),
                                        ptr(self.rng), ptr(err), ptr(self.colpart), self.block_n, cur_stream()),
              "mpit_umma_igemm")


@dataclass
class ResidentLaunch:
    """mpit_umma_conv_resident: one CTA per image, input planes resident in shared memory, weights streamed."""
    tmap_in: object
    tmap_w: object
    geom: object          # c_int * 15
    scale: float
    bias: Optional[torch.Tensor]
    out: torch.Tensor
    aux: Optional[torch.Tensor]
    aux_idx: Optional[torch.Tensor]
    block_n: int
    colpart: Optional[torch.Tensor] = None

    def run(self, err: torch.Tensor, dropout: Optional[bool] = None):
        check(kernels().mpit_umma_conv_resident(self.tmap_in, self.tmap_w, self.geom, self.scale, ptr(self.bias),
                                                ptr(self.out), ptr(self.aux), ptr(self.aux_idx), ptr(err),
                                                ptr(self.colpart), self.block_n, cur_stream()),
              "mpit_umma_conv_resident")

    @property
    def tiles(self) -> int:
        return self.geom[9] * self.geom[8]      # B * ntiles rows of colpart


def resident_geom(*, RW, RH, x_org, y_org, cb, KH, KW, sgn, ntiles, B, OH, OW, N, mode, relu=0):
    return (ctypes.c_int * 15)(RW, RH, x_org, y_org, cb, KH, KW, sgn, ntiles, B, OH, OW, N, mode, relu)


def igemm_geom(*, nsub, tw, th, tn, tiles_w, tiles_h, kw_it, cb, num_kb, sgn, B, OH, OW, N, mode, relu=0,
               dropout=0):
    assert nsub * tw * th * tn == 128
    return (ctypes.c_int * 17)(nsub, tw, th, tn, tiles_w, tiles_h, kw_it, cb, num_kb, sgn, B, OH, OW, N, mode,
                               int(relu), int(dropout))


EPI_STORE, EPI_POOL, EPI_MASK, EPI_EXPAND = 0, 1, 2, 3


def nhwc_tmap(t: torch.Tensor, B: int, H: int, W: int, C: int, box_w: int, box_h: int, box_n: int,
              swizzle: int = 1):
    """4-D map over an NHWC fp32 tensor, box = (32 channels, box_w, box_h, box_n)."""
    return make_tmap(t.data_ptr(), [C, W, H, B], [C * 4, W * C * 4, H * W * C * 4], [32, box_w, box_h, box_n],
                     swizzle)


def weight_tmap(t: torch.Tensor, rows: int, K: int, block_n: int):
    """2-D map over a [rows][K] K-major operand matrix, box = (32, block_n)."""
    return make_tmap(t.data_ptr(), [K, rows], [K * 4], [32, block_n])


def _pick_block(n: int, prefer=(128, 64, 32)) -> int:
    for b in prefer:
        if n % b == 0:
            return b
    raise ValueError(f"N={n} is not a multiple of 32")


class CifarUmmaPlan:
    """Launch plan of the Conv5-pool / Conv5-pool / Conv3 / FC / FC network on the tensor cores
    (forward, data gradients, weight gradients) for any channel counts the layer table allows."""

    def __init__(self, eng, wgrad: str = "auto"):
        self.eng = eng
        B = eng.B
        assert B % 2 == 0, "the 128-row tiles pair two 8x8 sub-tiles: batch must be even"
        L1, L2, L3, L4, L5 = eng.layers
        c1, c2, c3, f_in, f1 = L1.CO, L2.CO, L3.CO, L4.C, L4.CO
        self.err = torch.zeros(1, dtype=torch.int32, device=eng.dev)
        act, dz, idx = eng.act, eng.dz, eng.idx
        w = lambda i: eng.w(i)

        def wd(i):
            L = eng.layers[i]
            return eng.wpk[L.d_off:L.d_off + L.C * L.k * L.k * L.CO]

        self.fwd: List[IgemmLaunch] = []
        # conv1: overlapping view of the pair-expanded NHWC8 image: one 128-byte row = 4 pixels x 8 floats
        # = kernel columns [0,1|1,2|2,3|3,4] x 4 channels, pixel stride 32 bytes
        ta = make_tmap(eng.x0.data_ptr(), [32, 32, 32, B], [32, 32 * 32, 32 * 32 * 32], [32, 8, 8, 1])
        bn = _pick_block(c1, (64, 32))
        self.fwd.append(IgemmLaunch(
            ta, weight_tmap(w(0), c1, 5 * 32, bn),
            igemm_geom(nsub=2, tw=8, th=8, tn=1, tiles_w=3, tiles_h=3, kw_it=1, cb=1, num_kb=5, sgn=1, B=B,
                       OH=24, OW=24, N=c1, mode=EPI_POOL, relu=1),
            1.0, eng.b(0), act[0], idx[0], None, None, eng.crop, None, bn))
        # conv2: 64-column tiles -> twice the CTAs (128 instead of 64 for c2 = 128).  At TF32 the tensor pipe needs 128
        # clocks per 128x128x8 MMA, so 64 CTAs were MMA-bound at 13.5 us; MPIT_CONV2_BN overrides for A/B runs.
        import os as _os2
        bn = _pick_block(c2, tuple(int(v) for v in _os2.environ.get("MPIT_CONV2_BN", "64,32").split(",")))
        self.fwd.append(IgemmLaunch(
            nhwc_tmap(act[0], B, 12, 12, c1, 8, 8, 1), weight_tmap(w(1), c2, 25 * c1, bn),
            igemm_geom(nsub=2, tw=8, th=8, tn=1, tiles_w=1, tiles_h=1, kw_it=5, cb=c1 // 32, num_kb=25 * c1 // 32,
                       sgn=1, B=B, OH=8, OW=8, N=c2, mode=EPI_POOL, relu=1),
            1.0, eng.b(1), act[1], idx[1], None, None, None, None, bn))
        # conv3
        self.fwd.append(IgemmLaunch(
            nhwc_tmap(act[1], B, 4, 4, c2, 2, 2, 32), weight_tmap(w(2), c3, 9 * c2, 32),
            igemm_geom(nsub=1, tw=2, th=2, tn=32, tiles_w=1, tiles_h=1, kw_it=3, cb=c2 // 32, num_kb=9 * c2 // 32,
                       sgn=1, B=B, OH=2, OW=2, N=c3, mode=EPI_STORE, relu=1),
            1.0, eng.b(2), act[2], None, None, None, None, None, 32))
        # fc1 (+ReLU +dropout)
        self.fwd.append(IgemmLaunch(
            nhwc_tmap(act[2], B, 1, 1, f_in, 1, 1, 128), weight_tmap(w(3), f1, f_in, 32),
            igemm_geom(nsub=1, tw=1, th=1, tn=128, tiles_w=1, tiles_h=1, kw_it=1, cb=f_in // 32, num_kb=f_in // 32,
                       sgn=1, B=B, OH=1, OW=1, N=f1, mode=EPI_STORE, relu=1, dropout=1),
            1.0, eng.b(3), act[3], None, None, None, None, eng.rng, 32))
        # fc2
        self.fwd.append(IgemmLaunch(
            nhwc_tmap(act[3], B, 1, 1, f1, 1, 1, 128), weight_tmap(w(4), 32, f1, 32),
            igemm_geom(nsub=1, tw=1, th=1, tn=128, tiles_w=1, tiles_h=1, kw_it=1, cb=f1 // 32, num_kb=f1 // 32,
                       sgn=1, B=B, OH=1, OW=1, N=32, mode=EPI_STORE),
            1.0, eng.b(4), act[4], None, None, None, None, None, 32))

        # ---- data gradients (layer i -> dz of layer i-1) ----
        self.dgrad = {}
        # fc2: dz5 [B,32] x Wd5 [f1][32] -> mask by a4 > 0, x2 (dropout scale) -> dz4
        self.dgrad[4] = IgemmLaunch(
            nhwc_tmap(dz[4], B, 1, 1, 32, 1, 1, 128), weight_tmap(wd(4), f1, 32, 32),
            igemm_geom(nsub=1, tw=1, th=1, tn=128, tiles_w=1, tiles_h=1, kw_it=1, cb=1, num_kb=1, sgn=-1, B=B,
                       OH=1, OW=1, N=f1, mode=EPI_MASK),
            2.0, None, dz[3], None, act[3], None, None, None, 32)
        # fc1: dz4 [B,f1] x Wd4 [f_in][f1] -> mask by a3 > 0 -> dz3
        self.dgrad[3] = IgemmLaunch(
            nhwc_tmap(dz[3], B, 1, 1, f1, 1, 1, 128), weight_tmap(wd(3), f_in, f1, 32),
            igemm_geom(nsub=1, tw=1, th=1, tn=128, tiles_w=1, tiles_h=1, kw_it=1, cb=f1 // 32, num_kb=f1 // 32,
                       sgn=-1, B=B, OH=1, OW=1, N=f_in, mode=EPI_MASK),
            1.0, None, dz[2], None, act[2], None, None, None, 32)
        # conv3: dz3 [B,2,2,c3] -> d(a2) over 4x4 -> un-pool with idx2, mask a2 > 0 -> dz2 [B,8,8,c2]
        bn = _pick_block(c2, (64, 32))
        self.dgrad[2] = IgemmLaunch(
            nhwc_tmap(dz[2], B, 2, 2, c3, 4, 4, 8), weight_tmap(wd(2), c2, 9 * c3, bn),
            igemm_geom(nsub=1, tw=4, th=4, tn=8, tiles_w=1, tiles_h=1, kw_it=3, cb=c3 // 32, num_kb=9 * c3 // 32,
                       sgn=-1, B=B, OH=4, OW=4, N=c2, mode=EPI_EXPAND),
            1.0, None, dz[1], None, act[1], idx[1], None, None, bn)
        # conv2: dz2 [B,8,8,c2] -> d(a1) over 12x12 -> un-pool with idx1, mask -> dz1 [B,24,24,c1]
        bn = _pick_block(c1, (64, 32))
        # tile shape (MPIT_DGRAD1_TILE = "nsub,tw,th,tn"): rows of a 128-row tile = tn images x th x tw pixels
        import os as _os3
        d1 = [int(v) for v in _os3.environ.get("MPIT_DGRAD1_TILE", "1,4,4,8").split(",")]
        self.dgrad[1] = IgemmLaunch(
            nhwc_tmap(dz[1], B, 8, 8, c2, d1[1], d1[2], d1[3]), weight_tmap(wd(1), c1, 25 * c2, bn),
            igemm_geom(nsub=d1[0], tw=d1[1], th=d1[2], tn=d1[3], tiles_w=12 // d1[1], tiles_h=12 // d1[2], kw_it=5,
                       cb=c2 // 32, num_kb=25 * c2 // 32, sgn=-1, B=B, OH=12, OW=12, N=c1, mode=EPI_EXPAND),
            1.0, None, dz[0], None, act[0], idx[0], None, None, bn)

        # Bias gradients ride on the data-gradient epilogues: dgrad[i] writes dz[i-1], and its per-tile column
        # sums (summed over tiles by the split-K reduce launch) are the bias gradient of layer i-1.  The
        # classifier's bias gradient comes from the loss kernel.  (tiles, columns seen by the kernel, channels)
        # Optional (MPIT_RESIDENT=1): conv2's data gradient with the image's dz resident in shared memory (25 taps = 25
        # descriptor offsets into the same planes) instead of 25 re-fetched windows.  Correct, and it removes 90 % of
        # the L2 -> SM traffic, but it is SLOWER on B200 (50.6 us vs 27.8 us, profiles/README.md): one image per CTA in
        # a 16-pixel padded row pitch computes 256 rows for 144 useful ones, and at TF32 the tensor pipe (64 clk per
        # 128x64x8 MMA) then needs 27 us where the streaming kernel needs 13.5 us.  The streaming launch is the
        # default; it also stays available as self.dgrad_stream[1].
        self.dgrad_stream = dict(self.dgrad)
        import os as _os
        RW = 16
        ntl = -(-(12 * RW) // 128)
        RH = -(-(ntl * 128 + 4 * RW + 4) // RW)
        if (RW * RH) % 8:
            RH += 1
        plane = RW * RH * 128
        if (_os.environ.get("MPIT_RESIDENT", "0") == "1" and c1 == 64 and c2 % 32 == 0
                and (c2 // 32) * plane + 2 * c1 * 128 + 2048 <= 220 * 1024 and ntl * c1 <= 512):
            self.dgrad[1] = ResidentLaunch(
                nhwc_tmap(dz[1], B, 8, 8, c2, RW, RH, 1), weight_tmap(wd(1), c1, 25 * c2, c1),
                resident_geom(RW=RW, RH=RH, x_org=-4, y_org=-4, cb=c2 // 32, KH=5, KW=5, sgn=-1, ntiles=ntl, B=B,
                              OH=12, OW=12, N=c1, mode=EPI_EXPAND),
                1.0, None, dz[0], act[0], idx[0], c1)

        # The classifier (fc2 + log-softmax + NLL and their backward down to dz of fc1) is one CUDA-core kernel
        # (csrc/kernels/head.cu) when its shape allows: it then also produces gb(4) and gb(3) directly.
        import os
        self.head = os.environ.get("MPIT_HEAD", "1") == "1" and f1 in (64, 128, 256, 512) and L5.CO <= 32
        self._head_scratch = eng.ops.classifier_head_scratch(f1, eng.dev) if self.head else None
        self.bias_parts = {}
        for i in ((3, 2, 1) if self.head else (4, 3, 2, 1)):
            g = self.dgrad[i].geom            # {nsub,tw,th,tn,tiles_w,tiles_h,...,B at [10],...,N at [13]}
            if isinstance(self.dgrad[i], ResidentLaunch):
                tiles, ncols = self.dgrad[i].tiles, g[12]
            else:
                tiles = -(-(-(-g[10] // g[3]) * g[4] * g[5]) // g[0])
                ncols = g[13]
            cp = torch.zeros(tiles * ncols, dtype=torch.float32, device=eng.dev)
            self.dgrad[i].colpart = cp
            self.bias_parts[i - 1] = (cp, tiles, ncols, eng.layers[i - 1].CO)

        for l in list(self.fwd) + list(self.dgrad.values()) + list(self.dgrad_stream.values()):
            if isinstance(l, IgemmLaunch):
                l.x3 = getattr(eng, "x3", 0)
        self.two_streams = os.environ.get("MPIT_BWD_STREAMS", "2") != "1"
        # One more fork exists but is off by default: reducing layers 1..4 on the side stream beside the last data
        # gradient.  Every fork/join turns a programmatic (PDL) edge into a full one, and on B200 the two effects
        # cancel (A/B, ms/step: PDL on + no fork 0.1619; PDL on + fork 0.1634; PDL off + fork 0.1661).
        self.split_reduce = os.environ.get("MPIT_SPLIT_REDUCE", "0") == "1"
        self._side = torch.cuda.Stream(device=eng.dev)
        self._ev = [torch.cuda.Event() for _ in range(5)]
        self.wgrad = None
        if wgrad != "ref":
            try:
                from .umma_wgrad_ops import CifarWgradPlan
                self.wgrad = CifarWgradPlan(eng, self.err, self.bias_parts)
            except ImportError:
                if wgrad == "umma":
                    raise

    # -- schedules -----------------------------------------------------------------------------
    def forward(self, train: bool, classifier: bool = True):
        for i, l in enumerate(self.fwd if classifier else self.fwd[:4]):
            l.run(self.err, dropout=(train if i == 3 else None))

    def head_loss(self, train: bool) -> bool:
        """fc2 + loss (+ backward to dz[3], gb(4), gb(3) when training) in one launch; False if not applicable."""
        if not self.head:
            return False
        eng = self.eng
        L4, L5 = eng.layers[3], eng.layers[4]
        return eng.ops.classifier_head(
            eng.act[3], eng.w(4), eng.b(4), eng.labels, eng.act[4], eng.dz[4] if train else None,
            eng.dz[3] if train else None, eng.loss_sum, eng.B, L4.CO, eng.ncls, L5.CO, scratch=self._head_scratch,
            gscale=1.0 / eng.B,
            in_scale=2.0 if L4.dropout else 1.0, confusion=eng.confusion,
            db_cls=eng.gb(4) if train else None, db_in=eng.gb(3) if train else None,
            rng_inc=eng.rng if train else None)

    def backward(self):
        """Backward sweep.  The data-gradient chain (dgrad 4 -> 3 -> 2 -> 1) is the critical path; the weight
        gradient of layer i only needs dz[i], so it runs on a side stream next to the chain (most of these
        kernels occupy a handful of SMs).  The two streams join before the one reduce launch.  Captured into
        a CUDA graph this becomes two parallel branches."""
        eng, ops, B = self.eng, self.eng.ops, self.eng.B
        if self.wgrad is None:
            for i in range(4, -1, -1):
                L = eng.layers[i]
                x_in = eng.act[i - 1] if i > 0 else eng.x0
                ops.conv_wgrad_ref(L.geom(B), eng.dz[i], x_in, eng.gw(i), crop=eng.crop if i == 0 else None)
                ops.colsum(eng.dz[i], eng.gb(i), B * L.out_hw * L.out_hw, L.CO)
                if i > 0:
                    self.dgrad[i].run(self.err)
            return
        main = torch.cuda.current_stream()
        skip = 4 if self.head else -1      # the head kernel already produced dz[3]
        if not self.two_streams:
            for i in range(4, -1, -1):
                self.wgrad.run(i)
                if i > 0 and i != skip:
                    self.dgrad[i].run(self.err)
            self.wgrad.reduce()
            return
        side = self._side
        for i in range(4, 0, -1):
            self._ev[i].record(main)            # dz[i] is final on the main stream
            side.wait_event(self._ev[i])
            with torch.cuda.stream(side):
                self.wgrad.run(i)
            if i != skip:
                self.dgrad[i].run(self.err)
        if self.split_reduce:
            with torch.cuda.stream(side):       # layers 1..4 are final on the side stream: reduce them there,
                self.wgrad.reduce("early")      # beside the last (longest) data gradient
        self.wgrad.run(0)                       # needs dz[0] from dgrad[1]: stays on the main stream
        self._ev[0].record(side)
        main.wait_event(self._ev[0])
        self.wgrad.reduce("late" if self.split_reduce else "all")

    def check(self):
        """Raise if any bounded pipeline wait inside a tcgen05 kernel expired."""
        e = int(self.err.item())
        if e:
            raise RuntimeError(f"umma pipeline wait timed out (code {e})")
