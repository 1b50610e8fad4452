"""Layer kernels: fp32 CUDA-core reference path (csrc/kernels/nn_ref.cu)."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

from .._native import kernels, check, ptr, cur_stream


@dataclass(frozen=True)
class ConvGeom:
    """Geometry of one valid (no padding, stride 1) convolution on NHWC activations.

    ``SH,SW`` is the stored image, ``H,W`` the (possibly cropped) window the conv sees,
    ``C`` the physical (padded) input channels, ``KWP`` the physical kernel width and ``CO``
    the physical output channels of the packed ``[CO][KH][KWP][C]`` weight.
    A Linear layer is the 1x1 case (H=W=KH=KW=1).
    """
    B: int
    SH: int
    SW: int
    H: int
    W: int
    C: int
    KH: int
    KW: int
    KWP: int
    OH: int
    OW: int
    CO: int

    def carray(self):
        return (ctypes.c_int * 12)(self.B, self.SH, self.SW, self.H, self.W, self.C, self.KH, self.KW,
                                   self.KWP, self.OH, self.OW, self.CO)

    @property
    def wsize(self):
        return self.CO * self.KH * self.KWP * self.C


def conv_fwd_ref(g: ConvGeom, x, w, bias, out, *, crop=None, pool_idx=None, relu=False, pool=False,
                 dropout=False, rng=None):
    check(kernels().mpit_conv_fwd_ref(g.carray(), ptr(x), ptr(w), ptr(bias), ptr(crop), ptr(out),
                                      ptr(pool_idx), int(relu), int(pool), int(dropout), ptr(rng),
                                      cur_stream()), "mpit_conv_fwd_ref")


def conv_dgrad_ref(g: ConvGeom, dz, w, dx):
    check(kernels().mpit_conv_dgrad_ref(g.carray(), ptr(dz), ptr(w), ptr(dx), cur_stream()),
          "mpit_conv_dgrad_ref")


def conv_wgrad_ref(g: ConvGeom, dz, x, dw, *, crop=None):
    check(kernels().mpit_conv_wgrad_ref(g.carray(), ptr(dz), ptr(x), ptr(crop), ptr(dw), cur_stream()),
          "mpit_conv_wgrad_ref")


def act_bwd(dout, out, dz, B, OH, OW, C, *, pool_idx=None, pool=False, scale=1.0):
    check(kernels().mpit_act_bwd(ptr(dout), ptr(out), ptr(pool_idx), ptr(dz), B, OH, OW, C, int(pool),
                                 scale, cur_stream()), "mpit_act_bwd")


def colsum(dz, db, rows, C):
    check(kernels().mpit_colsum(ptr(dz), ptr(db), rows, C, cur_stream()), "mpit_colsum")


def prep_u8(src, dst, B, SH, SW, cin=3, labels=None, labels_dst=None, crop=None, crop_dst=None):
    """uint8 NCHW -> normalised pair-expanded NHWC8; optionally also moves the batch's labels / crop offsets."""
    check(kernels().mpit_prep_u8(ptr(src), ptr(dst), B, SH, SW, cin, ptr(labels), ptr(labels_dst), ptr(crop),
                                 ptr(crop_dst), cur_stream()), "mpit_prep_u8")


def nll_fwd_bwd(logits, labels, dlogits, loss_sum, B, ncls, ld, *, gscale, confusion=None, rng_inc=None,
                accumulate=False, dbias=None):
    """``loss_sum`` is overwritten with the batch's summed NLL (added to when ``accumulate``); ``dbias``
    (optional) receives the column sums of ``dlogits``: the classifier's bias gradient."""
    check(kernels().mpit_nll_fwd_bwd(ptr(logits), ptr(labels), ptr(dlogits), ptr(loss_sum), ptr(confusion),
                                     B, ncls, ld, gscale, ptr(rng_inc), int(accumulate), ptr(dbias), cur_stream()),
          "mpit_nll_fwd_bwd")


def classifier_head_scratch(K: int, device) -> "torch.Tensor":
    """Zero-filled workspace for :func:`classifier_head` (allocate once, reuse for every call)."""
    import torch
    return torch.zeros(kernels().mpit_classifier_head_scratch_floats(K), dtype=torch.float32, device=device)


def classifier_head(x, w, bias, labels, logits, dlogits, dx, loss_sum, B, K, ncls, ld, *, scratch, gscale, in_scale=1.0,
                    confusion=None, db_cls=None, db_in=None, rng_inc=None, accumulate=False) -> bool:
    """Fused Linear(K->ncls) + LogSoftMax + NLL forward/backward (csrc/kernels/head.cu).  Returns False when the
    shape is outside the kernel's limits (the caller then uses the GEMM + loss kernels)."""
    code = kernels().mpit_classifier_head(ptr(x), ptr(w), ptr(bias), ptr(labels), ptr(logits), ptr(dlogits), ptr(dx),
                                          ptr(loss_sum), ptr(confusion), ptr(db_cls), ptr(db_in), ptr(rng_inc),
                                          ptr(scratch), B, K, ncls, ld, gscale, in_scale, int(accumulate),
                                          cur_stream())
    if code == -2:      # MPIT_E_UNSUPPORTED
        return False
    check(code, "mpit_classifier_head")
    return True
