"""Launch plan of the tcgen05 weight-gradient kernels (csrc/kernels/umma_wgrad.cu) for the CIFAR CNN."""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass
from typing import Optional

import torch

from .._native import kernels, check, ptr, cur_stream
from .umma_ops import make_tmap, nhwc_tmap as _nhwc_tmap

MN_SWIZZLE = 2   # CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B: the only layout tf32 MN-major operands support


def nhwc_tmap(*a):
    return _nhwc_tmap(*a, swizzle=MN_SWIZZLE)


@dataclass
class WgradLaunch:
    tmap_x: object
    tmap_dz: object
    geom: object
    crop: Optional[torch.Tensor]
    block_n: int
    nsplit: int
    m_pad: int
    K: int
    N: int
    gw: torch.Tensor
    part_off: int


class CifarWgradPlan:
    def __init__(self, eng, err: torch.Tensor, bias_parts=None):
        self.eng, self.err = eng, err
        B = eng.B
        act, dz = eng.act, eng.dz
        G = lambda *a: (ctypes.c_int * 11)(*a)
        ceil = lambda a, b: (a + b - 1) // b
        x0map = make_tmap(eng.x0.data_ptr(), [32, 32, 32, B], [32, 32 * 32, 32 * 32 * 32], [32, 8, 8, 1], MN_SWIZZLE)
        L1, L2, L3, L4, L5 = eng.layers
        c1, c2, c3, f_in, f1 = L1.CO, L2.CO, L3.CO, L4.C, L4.CO
        pick = lambda n: 128 if n % 128 == 0 else (64 if n % 64 == 0 else 32)
        specs = [
            # tmap_x, tmap_dz, (tw,th,tn,tiles_w,tiles_h,kw_it,cb,total_chunks,num_kb,B,N), crop, block_n, nsplit
            (x0map, nhwc_tmap(dz[0], B, 24, 24, c1, 8, 8, 1), (8, 8, 1, 3, 3, 1, 1, 5, B * 9, B, c1), eng.crop,
             min(64, pick(c1)), min(int(os.environ.get("MPIT_WGRAD1_NSPLIT", "64")), B * 9)),
            (nhwc_tmap(act[0], B, 12, 12, c1, 8, 8, 1), nhwc_tmap(dz[1], B, 8, 8, c2, 8, 8, 1),
             (8, 8, 1, 1, 1, 5, c1 // 32, 25 * c1 // 32, B, B, c2), None, pick(c2), min(8, B)),
            (nhwc_tmap(act[1], B, 4, 4, c2, 2, 2, 16), nhwc_tmap(dz[2], B, 2, 2, c3, 2, 2, 16),
             (2, 2, 16, 1, 1, 3, c2 // 32, 9 * c2 // 32, ceil(B, 16), B, c3), None, min(64, pick(c3)), ceil(B, 16)),
            (nhwc_tmap(act[2], B, 1, 1, f_in, 1, 1, 64), nhwc_tmap(dz[3], B, 1, 1, f1, 1, 1, 64),
             (1, 1, 64, 1, 1, 1, f_in // 32, f_in // 32, ceil(B, 64), B, f1), None, pick(f1), ceil(B, 64)),
            (nhwc_tmap(act[3], B, 1, 1, f1, 1, 1, 64), nhwc_tmap(dz[4], B, 1, 1, 32, 1, 1, 64),
             (1, 1, 64, 1, 1, 1, f1 // 32, f1 // 32, ceil(B, 64), B, 32), None, 32, ceil(B, 64)),
        ]
        self.launches = []
        off = 0
        for i, (tx, tz, g, crop, bn, ns) in enumerate(specs):
            chunks, N = g[7], g[10]
            m_pad = ceil(chunks, 4) * 128
            self.launches.append(WgradLaunch(tx, tz, G(*g), crop, bn, ns, m_pad, chunks * 32, N, eng.gw(i), off))
            off += ns * m_pad * N
        # split-K scratch workspace shared by all layers
        self.part = torch.empty(off, dtype=torch.float32, device=eng.dev)
        ls = self.launches
        # reduce segments: split-K partials of every layer's weight gradient ...
        wsegs = {i: (self.part.data_ptr() + 4 * l.part_off, l.gw.data_ptr(), l.nsplit, l.m_pad, l.K, l.N)
                 for i, l in enumerate(ls)}
        # ... and the bias-gradient partial rows written by the data-gradient epilogues: T tile rows of
        # G*C columns -> {nsplit = T*G, m_pad = 1, K = 1, N = C}
        self.bias_from_epilogue = bias_parts is not None
        bsegs = {}
        if bias_parts:
            for li, (cp, tiles, ncols, C) in sorted(bias_parts.items()):
                assert ncols % C == 0
                bsegs[li] = (cp.data_ptr(), eng.gb(li).data_ptr(), tiles * (ncols // C), 1, 1, C)

        def pack(segs):
            n = len(segs)
            return (n, (ctypes.c_ulonglong * n)(*[g[0] for g in segs]), (ctypes.c_ulonglong * n)(*[g[1] for g in segs]),
                    (ctypes.c_int * n)(*[g[2] for g in segs]), (ctypes.c_int * n)(*[g[3] for g in segs]),
                    (ctypes.c_int * n)(*[g[4] for g in segs]), (ctypes.c_int * n)(*[g[5] for g in segs]))

        # "all": one launch at the end.  "early"/"late": everything that is final once the layer-1 weight gradient
        # is done (layers 1..4 and the bias rows of the data gradients that ran before it) / the first layer.
        # with the optimizer epilogue every parameter has to pass through the reduction: bias gradients that come from
        # elsewhere (the classifier head's loss kernel) ride along as one-partial segments that read and rewrite gb
        extra = [(eng.gb(li).data_ptr(), eng.gb(li).data_ptr(), 1, 1, 1, eng.layers[li].CO)
                 for li in range(len(ls)) if li not in bsegs]
        self._red_opt = pack(list(wsegs.values()) + list(bsegs.values()) + extra) if bias_parts else None
        self._red = {"all": pack(list(wsegs.values()) + list(bsegs.values())),
                     "early": pack([wsegs[i] for i in range(1, len(ls))] + [bsegs[i] for i in sorted(bsegs) if i >= 1]),
                     "late": pack([wsegs[0]] + ([bsegs[0]] if 0 in bsegs else []))}
        L = eng.layers
        n = len(ls)
        self._cs = ((ctypes.c_ulonglong * n)(*[dz[i].data_ptr() for i in range(n)]),
                    (ctypes.c_ulonglong * n)(*[eng.gb(i).data_ptr() for i in range(n)]),
                    (ctypes.c_longlong * n)(*[B * L[i].out_hw * L[i].out_hw for i in range(n)]),
                    (ctypes.c_int * n)(*[L[i].CO for i in range(n)]))

    def run(self, i: int):
        l = self.launches[i]
        k = kernels()
        k.mpit_set_tf32x3(getattr(self.eng, "x3", 0))
        part = self.part[l.part_off:]
        check(k.mpit_umma_wgrad(l.tmap_x, l.tmap_dz, l.geom, ptr(part), ptr(l.crop), ptr(self.err), l.block_n,
                                l.nsplit, cur_stream()), "mpit_umma_wgrad")

    def reduce(self, which: str = "all"):
        """ONE launch sums split-K partials of weight gradients and the bias-gradient partial rows of the
        data-gradient epilogues.  ``which``: "all", or "early" (layers 1..4, runs beside the last data gradient)
        followed by "late" (first layer)."""
        k = kernels()
        o = getattr(self.eng, "opt_epi", None)
        if o is not None and which == "all" and self._red_opt is not None:
            # the reduction finishes the step: apply + next look-ahead + operand refresh (no boundary kernel next step)
            e, r = self.eng, self._red_opt
            check(k.mpit_multi_wgrad_reduce_opt(r[0], r[1], r[2], r[3], r[4], r[5], r[6], ptr(o["x"]), ptr(o["v"]),
                                                ptr(o["d"]), ptr(e.pk2flat), ptr(e.wpk), ptr(e.wmap_d), ptr(e.gpk),
                                                o["clr"], o["l2wd"], o["mom_next"], int(o["use_mom"]), int(o["has_d"]),
                                                int(o["la_in"]), cur_stream()), "mpit_multi_wgrad_reduce_opt")
            return
        r = self._red[which]
        check(k.mpit_multi_wgrad_reduce(r[0], r[1], r[2], r[3], r[4], r[5], r[6], cur_stream()),
              "mpit_multi_wgrad_reduce")
        if not self.bias_from_epilogue and which in ("all", "late"):
            n = len(self.launches)
            check(k.mpit_multi_colsum(n, self._cs[0], self._cs[1], self._cs[2], self._cs[3], cur_stream()),
                  "mpit_multi_colsum")
