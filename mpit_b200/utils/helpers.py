"""Small helpers of the BiCNN workload: time stamps, tensor descriptions, layer-size arithmetic, dict utilities and
deep clones with a dtype change.

Reference: /root/reference/BiCNN/utils.lua:4-254 (global Lua functions ``sleep, millisleep, tstamp, tsize, tdims,
tinfo, tdinfo, tflat, lnarrow, conv_size, pool_size, pad_size, count_list, sorted_keys, table_acc, table_string,
clone_table, clone_module, clone_object``).  None of them is on a hot path there or here.  Differences worth knowing:
``tsize`` handles any rank (the reference stops at six, utils.lua:20-35); ``tdinfo`` in the reference reads four
undefined globals (``weights, cap_conv, cap_linear, cap_bias``, utils.lua:58-96) -- here they are arguments and the
function is also exported under the name that says what it does, ``cap_weight_norms``; the clone helpers take a torch
dtype instead of a Torch7 tensor type name.
"""
from __future__ import annotations

import copy
import sys
import time
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple

import torch


def sleep(n: float) -> None:
    time.sleep(float(n))


def millisleep(n: float) -> None:
    time.sleep(float(n) / 1000.0)


def tstamp(t: Optional[float] = None) -> str:
    """``YYYYMMDD.HHMMSS`` of ``t`` (seconds since the epoch) or of now (utils.lua:14-18)."""
    return time.strftime("%Y%m%d.%H%M%S", time.localtime(t))


def tsize(size: Optional[Sequence[int]]) -> str:
    return "(no tensor)" if size is None else "x".join(str(int(s)) for s in size)


def tdims(t: Optional[torch.Tensor]) -> str:
    return "(no tensor)" if t is None else tsize(t.shape)


def tinfo(t) -> str:
    """One-line summary: shape, min, mean, max, dtype (utils.lua:45-56)."""
    if t is None:
        return "(no tensor)"
    if not isinstance(t, torch.Tensor):
        return str(t)
    if t.numel() == 0:
        return "(empty tensor)"
    f = t.float() if not t.is_floating_point() else t
    return f"({tdims(t)} min: {float(t.min()):f} mean: {float(f.mean()):f} max: {float(t.max()):f} type: {t.dtype})"


def cap_weight_norms(weights: Iterable[torch.Tensor], cap_conv: float = 0.0, cap_linear: float = 0.0,
                     cap_bias: float = 0.0, conv_dim: int = 0, lin_dim: int = 0) -> None:
    """Max-norm constraint, in place: every output feature of a 4-d (convolution) or 2-d (linear) weight whose L2 norm
    exceeds the cap is rescaled onto it; biases are clamped to ``[-cap_bias, cap_bias]``.  A cap of 0 disables that
    group (the body of ``tdinfo``, utils.lua:58-96)."""
    for w in weights:
        if w.dim() == 4 and cap_conv > 0:
            _cap_features(w, conv_dim, cap_conv)
        elif w.dim() == 2 and cap_linear > 0:
            _cap_features(w, lin_dim, cap_linear)
        elif w.dim() == 1 and cap_bias > 0:
            w.clamp_(-cap_bias, cap_bias)


tdinfo = cap_weight_norms


def _cap_features(w: torch.Tensor, dim: int, cap: float) -> None:
    others = [d for d in range(w.dim()) if d != dim]
    norms = torch.linalg.vector_norm(w, dim=others, keepdim=True)
    w.mul_(torch.where(norms > cap, cap / norms, torch.ones_like(norms)))


def tflat(t: Optional[torch.Tensor], stream=sys.stdout) -> None:
    """Print a 1-d tensor on one line (utils.lua:98-107)."""
    stream.write("(no tensor)" if t is None else "".join(f"{v} " for v in t.tolist()))
    stream.write("\n")


def lnarrow(l: Sequence, s: int, o: int) -> List:
    """``s + 1`` items of ``l`` starting at 1-based offset ``o`` -- the reference's loop runs ``o .. o+s`` inclusive
    (utils.lua:110-117)."""
    return list(l[o - 1:o + s])


def _size_line(kind: str, text: str, w1: float, h1: float, stream) -> Tuple[float, float]:
    stream.write(text + "\n")
    if w1 != int(w1) or h1 != int(h1):
        stream.write(f"warning: expected integers in {w1}x{h1}\n")
    return w1, h1


def conv_size(w, h, kw, kh, sw, sh, feat1, feat2, stream=sys.stdout) -> Tuple[float, float]:
    """Output width / height of a valid convolution, printed like the reference (utils.lua:120-129)."""
    w1, h1 = (w - (kw - sw)) / sw, (h - (kh - sh)) / sh
    return _size_line("conv", f"convolution\t{feat1}x{w}x{h}\t* {kw}x{kh} (stride {sw}x{sh})\t-> {feat2}x{w1:g}x{h1:g}",
                      w1, h1, stream)


def pool_size(w, h, kw, kh, sw, sh, stream=sys.stdout) -> Tuple[float, float]:
    w1, h1 = (w - (kw - sw)) / sw, (h - (kh - sh)) / sh
    return _size_line("pool", f"pooling   \t{w}x{h}   \t/ {kw}x{kh} (stride {sw}x{sh})\t-> {w1:g}x{h1:g}", w1, h1, stream)


def pad_size(w, h, pw, ph, stream=sys.stdout) -> Tuple[int, int]:
    w1, h1 = w + 2 * pw, h + 2 * ph
    stream.write(f"padding   \t{w}x{h}      \t* {pw}x{ph}x{pw}x{ph}         \t-> {w1}x{h1}\n")
    return w1, h1


def count_list(l) -> int:
    return sum(1 for _ in (l.items() if isinstance(l, dict) else l))


def sorted_keys(t: Dict) -> List:
    return sorted(t)


def table_acc(t1: Dict, t2: Dict) -> None:
    """Accumulate ``t2`` into ``t1``: numbers add, everything else is overwritten (utils.lua:166-174)."""
    for k, v in t2.items():
        if isinstance(t1.get(k), (int, float)) and not isinstance(t1.get(k), bool):
            t1[k] = t1[k] + v
        else:
            t1[k] = v


def table_string(tab: Dict, factor: float = 1.0, fmt: str = " %s: %.3f") -> str:
    """Sorted ``key value*factor`` pairs formatted with ``fmt`` (a %-format taking key and number); nested dicts in
    braces (utils.lua:178-193)."""
    s = ""
    for k in sorted_keys(tab):
        p = tab[k]
        if p is None or p is False:
            continue
        s += f" {k} {{{table_string(p, factor, fmt)} }}" if isinstance(p, dict) else fmt % (k, p * factor)
    return s


def clone_table(table: Dict, dtype: Optional[torch.dtype] = None, target: Optional[Dict] = None) -> Dict:
    """Deep copy of a (nested) dict; tensors are copied and converted to ``dtype`` when given (utils.lua:196-214)."""
    if table is None:
        raise ValueError("trying to clone a nil table")
    target = {} if target is None else target
    for k, v in table.items():
        if isinstance(v, torch.Tensor):
            target[k] = v.detach().clone().to(dtype) if dtype is not None and v.is_floating_point() else v.detach().clone()
        elif isinstance(v, dict):
            target[k] = clone_table(v, dtype)
        else:
            target[k] = v
    return target


def clone_module(module: torch.nn.Module, dtype: Optional[torch.dtype] = None,
                 exclude: Sequence[str] = ()) -> torch.nn.Module:
    """Deep copy of a module (sub-modules included) with its floating-point tensors converted to ``dtype``; tensors
    whose attribute name is listed in ``exclude`` are dropped from the copy, e.g. ``("gradInput", "output")`` in the
    reference's use (utils.lua:216-246)."""
    clone = copy.deepcopy(module)
    if dtype is not None:
        clone = clone.to(dtype)
    for m in clone.modules():
        for name in exclude:
            if name in m._parameters:
                m._parameters[name] = None
            elif name in m._buffers:
                m._buffers[name] = None
            elif isinstance(m.__dict__.get(name), torch.Tensor):
                del m.__dict__[name]
    return clone


def clone_object(o: Any) -> Any:
    """A new object of the same class whose attributes are ``clone_table`` copies (utils.lua:248-254)."""
    c = copy.copy(o)
    if hasattr(o, "__dict__"):
        c.__dict__ = clone_table(o.__dict__)
    return c
