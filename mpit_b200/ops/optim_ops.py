"""Optimizer kernels (csrc/kernels/optim.cu)."""
from __future__ import annotations

import ctypes
import math

from .._native import kernels, check, ptr, cur_stream

RULES = {"add": 0, "rmsprop": 1, "adam": 2, "adamax": 3, "adagrad": 4, "adadelta": 5}


def _n(t):
    return int(t.numel())


def _nmaps(wmap, n):
    """Number of stacked index tables in ``wmap`` ([nmaps * n] int32)."""
    return 0 if wmap is None else max(1, int(wmap.numel()) // max(n, 1))


def local_step(x, v=None, g=None, gmap=None, d=None, wpk=None, wmap=None, *, clr=0.0, l2wd=0.0,
               mom_next=0.0, do_apply=True, use_mom=False, do_lookahead=False):
    """Fused Nesterov apply/lookahead/pack pass (asyncsgd/optim-msgd.lua:20-39)."""
    check(kernels().mpit_local_step(ptr(x), ptr(v), ptr(g), ptr(gmap), ptr(d), ptr(wpk), ptr(wmap),
                                    _nmaps(wmap, _n(x)), _n(x), clr, l2wd, mom_next, int(do_apply), int(use_mom),
                                    int(do_lookahead), cur_stream()), "mpit_local_step")


def rule_hyper(rule: str, t: int, **kw) -> list:
    """Host-side scalar preparation for a rule at (1-based) update count ``t``.

    Mirrors the scalar arithmetic of BiCNN/pserver.lua:123-195 (server) and
    BiCNN/optim-*-single.lua (single); ``step_div`` reproduces the server's
    ``floor(t/stepDivAdam)+1`` bias-correction exponent (pserver.lua:151-152).
    Returns the six floats {a,b,c,d,eps,step} consumed by the kernel.
    """
    if rule == "add":
        return [0.0] * 6
    if rule == "rmsprop":
        decay, lr, mom, eps = kw["decay"], kw["lr"], kw["momentum"], kw["epsilon"]
        return [decay, 1 - decay, mom, 0.0, eps, -lr]
    if rule == "adam":
        b1, b2, lr, eps = kw["beta1"], kw["beta2"], kw["lr"], kw["epsilon"]
        sd = kw.get("step_div", 0)
        k = (t // sd + 1) if sd else t
        lr_t = lr * math.sqrt(1 - b2 ** k) / (1 - b1 ** k)
        return [b1, 1 - b1, b2, 1 - b2, eps, -lr_t]
    if rule == "adamax":
        b1, b2, lr, eps = kw["beta1"], kw["beta2"], kw["lr"], kw["epsilon"]
        lr_t = lr / (1 - b1 ** t)
        return [b1, 1 - b1, b2, 0.0, eps, -lr_t]
    if rule == "adagrad":
        lr, lrd, eps = kw["lr"], kw.get("lrd", 0.0), kw["epsilon"]
        clr = lr / (1 + (t - 1) * lrd)  # pversion counts completed updates
        return [0.0, 0.0, 0.0, 0.0, eps, -clr]
    if rule == "adadelta":
        rho, eps, lr = kw.get("rho", 0.9), kw.get("epsilon", 1e-6), kw["lr"]
        return [rho, 1 - rho, 0.0, 0.0, eps, -lr]
    raise ValueError(rule)


def rule_step(rule: str, p, g, hyper6, s0=None, s1=None, s2=None, gmap=None, wpk=None, wmap=None, *,
              l1=0.0, l2=0.0, clip=0.0, gscale=1.0):
    h = (ctypes.c_float * 6)(*hyper6)
    check(kernels().mpit_rule_step(RULES[rule], ptr(p), ptr(g), ptr(gmap), ptr(s0), ptr(s1), ptr(s2),
                                   ptr(wpk), ptr(wmap), _nmaps(wmap, _n(p)), _n(p), h, l1, l2, clip, gscale,
                                   cur_stream()),
          "mpit_rule_step")


def downpour_accum(x, acc, g, gmap=None, wpk=None, wmap=None, *, lr, l2wd=0.0, move_local=False):
    check(kernels().mpit_downpour_accum(ptr(x), ptr(acc), ptr(g), ptr(gmap), ptr(wpk), ptr(wmap),
                                        _nmaps(wmap, _n(x)), _n(x), -lr, l2wd, int(move_local), cur_stream()), "mpit_downpour_accum")


def grad_reg_clip(g, p, *, l1=0.0, l2=0.0, clip=0.0):
    check(kernels().mpit_grad_reg_clip(ptr(g), ptr(p), _n(g), l1, l2, clip, cur_stream()),
          "mpit_grad_reg_clip")


def map_copy(dst, src, map_, scatter=False):
    check(kernels().mpit_map_copy(ptr(dst), ptr(src), ptr(map_), int(map_.numel()), int(scatter),
                                  cur_stream()), "mpit_map_copy")
