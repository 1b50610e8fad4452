"""Generalised active-target synchronisation for host windows: post / start / complete / wait.

/root/reference/mpifuncs.c:2330,2434,2468,2493 wraps ``Win_post``, ``Win_start``, ``Win_complete``, ``Win_wait`` (and
``Win_test``, ``Win_get_group``) next to the fence calls ("One-sided 14/14"); the tree calls none
of them.  ``extras.py`` has the fence epochs; this module adds the pairwise epochs, in which only the processes named
in the two groups synchronise:

    target:  Win_post(origins)  ...............................  Win_wait()   (or Win_test() until True)
    origin:  Win_start(targets)  Put / Get / Accumulate ...  Win_complete()

As with fences, ``Put / Get / Accumulate`` are queued at the origin and travel at ``Win_complete``; the target applies
them in ``Win_wait``, in the order each origin issued them.  Groups hold WORLD ranks (``Comm_group`` /
``Group_incl``); the window remembers its communicator's group from ``Win_create``.  The device-memory counterpart
(NVLink ``st`` / ``ld`` / ``red.add`` with flag synchronisation) is ``mpit_b200.parallel.DeviceWindow``.
"""
from __future__ import annotations

from typing import List

import numpy as np
import torch

_TAG_POST, _TAG_DESC, _TAG_DATA, _TAG_GET = 32210, 32211, 32212, 32213
ERR_RMA_SYNC, ERR_GROUP = 50, 9
_KINDS = {"put": 0, "acc": 1, "get": 2}


def _m():
    from .. import msg
    return msg


def _err(code, what):
    return _m().MPIError(code, what)


def _comm_ranks(win, group) -> List[int]:
    """Ranks in the window's communicator of the world ranks in ``group``."""
    members = win.group.ranks
    try:
        return [members.index(r) for r in group.ranks]
    except ValueError:
        raise _err(ERR_GROUP, "the group names a process outside the window's communicator") from None


def Win_get_group(win):
    """The group of the communicator the window was created over."""
    return _m().Group(win.group.ranks)


def Win_post(group, assertion: int, win) -> None:
    """Target side: open an exposure epoch for the origins in ``group``."""
    m = _m()
    if getattr(win, "exposed_to", None):
        raise _err(ERR_RMA_SYNC, "Win_post: an exposure epoch is already open")
    win.exposed_to = _comm_ranks(win, group)
    for o in win.exposed_to:
        m.Send(None, 0, m.BYTE, o, _TAG_POST, win.comm)


def Win_start(group, assertion: int, win) -> None:
    """Origin side: open an access epoch to the targets in ``group``; returns once each has posted."""
    m = _m()
    if getattr(win, "access_to", None):
        raise _err(ERR_RMA_SYNC, "Win_start: an access epoch is already open")
    win.access_to = _comm_ranks(win, group)
    for t in win.access_to:
        m.Recv(None, 0, m.BYTE, t, _TAG_POST, win.comm)


def Win_complete(win) -> None:
    """Origin side: ship the queued operations to their targets and collect the results of the Gets."""
    m = _m()
    targets = getattr(win, "access_to", None)
    if targets is None:
        raise _err(ERR_RMA_SYNC, "Win_complete without Win_start")
    stray = [op for op in win.ops if op[1] not in targets]
    if stray:
        raise _err(ERR_RMA_SYNC, f"an operation targets rank {stray[0][1]}, which is outside the access group")
    for t in targets:
        mine = [op for op in win.ops if op[1] == t]
        desc = torch.tensor([len(mine)] + [v for o in mine for v in
                                           (_KINDS[o[0]], o[2], o[3], o[4], -1 if o[6] is None else o[6])],
                            dtype=torch.int64)
        m.Send(desc, desc.numel(), m.LONG, t, _TAG_DESC, win.comm)
        for o in mine:
            if o[0] != "get":
                m.Send(o[5], o[3], o[4], t, _TAG_DATA, win.comm)
    # every descriptor is on its way (sends are buffered) before the first Get result is awaited, so two origins
    # that list their targets in different orders cannot block each other
    for t in targets:
        for o in win.ops:
            if o[1] == t and o[0] == "get":
                m.Recv(o[5][:o[3]], o[3], o[4], t, _TAG_GET, win.comm)
    win.ops = []
    win.access_to = None


def _serve(win, origin: int) -> None:
    m = _m()
    ext = m.extras
    st = m.Status()
    m.Probe(origin, _TAG_DESC, win.comm, st)
    n = m.Get_count(st, m.LONG)
    desc = torch.empty(n, dtype=torch.int64)
    m.Recv(desc, n, m.LONG, origin, _TAG_DESC, win.comm)
    ops = desc[1:].view(-1, 5).tolist()
    is_t = isinstance(win.base, torch.Tensor)
    for kind, disp, cnt, dtp, op in ops:
        if kind == 2:
            continue
        tmp = torch.empty(cnt, dtype=win.base.dtype) if is_t else np.empty(cnt, win.base.dtype)
        m.Recv(tmp, cnt, dtp, origin, _TAG_DATA, win.comm)
        if kind == 0:
            win.base[disp:disp + cnt] = tmp
        else:
            ext._apply_op(win.base[disp:disp + cnt], tmp, op)
    for kind, disp, cnt, dtp, op in ops:
        if kind == 2:
            m.Send(win.base[disp:disp + cnt], cnt, dtp, origin, _TAG_GET, win.comm)


def Win_wait(win) -> None:
    """Target side: returns when every origin of the exposure epoch has called ``Win_complete``."""
    origins = getattr(win, "exposed_to", None)
    if origins is None:
        raise _err(ERR_RMA_SYNC, "Win_wait without Win_post")
    for o in origins:
        _serve(win, o)
    win.exposed_to = None


def Win_test(win) -> bool:
    """Non-blocking ``Win_wait``: True (epoch closed) once every origin's completion message has arrived."""
    m = _m()
    origins = getattr(win, "exposed_to", None)
    if origins is None:
        raise _err(ERR_RMA_SYNC, "Win_test without Win_post")
    if not all(m.Iprobe(o, _TAG_DESC, win.comm) for o in origins):
        return False
    Win_wait(win)
    return True


__all__ = [_n for _n in dir() if _n[0].isupper() and _n not in ("Any", "Callable", "Dict", "List", "Optional", "Sequence",
                                                                 "Tuple")]
