"""Host windows beyond fences: post / start / complete / wait epochs, and passive-target lock / unlock.

/root/reference/mpifuncs.c:2330,2434,2468,2493 wraps ``Win_post``, ``Win_start``, ``Win_complete``, ``Win_wait`` (and
``Win_test``, ``Win_get_group``) next to the fence calls ("One-sided 14/14"); the tree calls none
of them.  ``extras.py`` has the fence epochs; this module adds the pairwise epochs, in which only the processes named
in the two groups synchronise:

    target:  Win_post(origins)  ...............................  Win_wait()   (or Win_test() until True)
    origin:  Win_start(targets)  Put / Get / Accumulate ...  Win_complete()

As with fences, ``Put / Get / Accumulate`` are queued at the origin and travel at ``Win_complete``; the target applies
them in ``Win_wait``, in the order each origin issued them.  Groups hold WORLD ranks (``Comm_group`` /
``Group_incl``); the window remembers its communicator's group from ``Win_create``.

Passive target (``Win_lock`` / ``Win_unlock``, /root/reference/mpifuncs.c:2424,2485): the target takes no part in the
epoch, so every window runs a small service thread on a communicator of its own (created in the collective
``Win_create``).  It grants shared / exclusive locks in arrival order and applies ``Put / Accumulate / Get`` requests of
the lock holders as they arrive -- inside a lock the operations are sent at once instead of being queued; ``Win_unlock``
returns when the target has applied everything the origin sent.  The native message layer is thread-safe, which is all
this needs.

The device-memory counterpart (NVLink ``st`` / ``ld`` / ``red.add`` with flag synchronisation) is
``mpit_b200.parallel.DeviceWindow``.
"""
from __future__ import annotations

import threading
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



# ---- passive target: lock / unlock served by a thread at the target ------------------------------------------------
LOCK_EXCLUSIVE, LOCK_SHARED = 234, 235
_TAG_REQ, _TAG_PAY, _TAG_REPLY = 32220, 32221, 32222
_K_LOCK, _K_UNLOCK, _K_PUT, _K_ACC, _K_GET, _K_QUIT = range(6)


class _PassiveServer(threading.Thread):
    """Serves one window's passive-target requests.  Requests are 6 x int64 {kind, disp, count, dtype, op, lock type};
    Put / Accumulate payloads follow on their own tag; grants, Get data and unlock acknowledgements go back on the
    reply tag.  Requests of one origin arrive in the order it sent them, so an UNLOCK is processed after every
    operation of that epoch."""

    def __init__(self, win, comm: int):
        super().__init__(daemon=True, name="mpit-win-server")
        self.win, self.comm = win, comm
        self.exclusive = None            # origin holding the exclusive lock
        self.shared = set()              # origins holding a shared lock
        self.waiting = []                # (origin, lock type) in arrival order

    def _grantable(self, kind: int) -> bool:
        if kind == LOCK_EXCLUSIVE:
            return self.exclusive is None and not self.shared
        return self.exclusive is None

    def _grant(self, origin: int, kind: int) -> None:
        m = _m()
        if kind == LOCK_EXCLUSIVE:
            self.exclusive = origin
        else:
            self.shared.add(origin)
        m.Send(None, 0, m.BYTE, origin, _TAG_REPLY, self.comm)

    def _drain_waiting(self) -> None:
        while self.waiting and self._grantable(self.waiting[0][1]):
            origin, kind = self.waiting.pop(0)
            self._grant(origin, kind)

    def run(self) -> None:
        m = _m()
        ext = m.extras
        win = self.win
        hdr = torch.zeros(6, dtype=torch.int64)
        st = m.Status()
        while True:
            m.Recv(hdr, 6, m.LONG, m.ANY_SOURCE, _TAG_REQ, self.comm, st)
            origin = st.source
            kind, disp, cnt, dtp, op, ltype = hdr.tolist()
            if kind == _K_QUIT:
                return
            if kind == _K_LOCK:
                if self._grantable(ltype) and not self.waiting:
                    self._grant(origin, ltype)
                else:
                    self.waiting.append((origin, ltype))
            elif kind == _K_UNLOCK:
                if self.exclusive == origin:
                    self.exclusive = None
                self.shared.discard(origin)
                m.Send(None, 0, m.BYTE, origin, _TAG_REPLY, self.comm)
                self._drain_waiting()
            elif kind in (_K_PUT, _K_ACC):
                is_t = isinstance(win.base, torch.Tensor)
                tmp = torch.empty(cnt, dtype=win.base.dtype) if is_t else np.empty(cnt, win.base.dtype)
                m.Recv(tmp, cnt, dtp, origin, _TAG_PAY, self.comm)
                if kind == _K_PUT:
                    win.base[disp:disp + cnt] = tmp
                else:
                    ext._apply_op(win.base[disp:disp + cnt], tmp, op)
            elif kind == _K_GET:
                m.Send(win.base[disp:disp + cnt], cnt, dtp, origin, _TAG_REPLY, self.comm)


def start_passive_server(win) -> None:
    """Called from the collective ``Win_create``."""
    m = _m()
    if not m.Initialized():
        win.server = None
        return
    win.server_comm = m.Comm_dup(win.comm)
    win.locked = {}
    win.server = _PassiveServer(win, win.server_comm)
    win.server.start()


def stop_passive_server(win) -> None:
    m = _m()
    srv = getattr(win, "server", None)
    if srv is None:
        return
    if getattr(win, "locked", None):
        raise _err(ERR_RMA_SYNC, "Win_free with a lock still held")
    m.Barrier(win.comm)                                      # nobody is still talking to anybody's server
    quit_ = torch.tensor([_K_QUIT, 0, 0, 0, 0, 0], dtype=torch.int64)
    m.Send(quit_, 6, m.LONG, m.Comm_rank(win.server_comm), _TAG_REQ, win.server_comm)
    srv.join()
    m.Comm_free(win.server_comm)
    win.server = None


def _request(win, target: int, kind: int, disp=0, cnt=0, dtp=0, op=-1, ltype=0) -> None:
    m = _m()
    m.Send(torch.tensor([kind, disp, cnt, dtp, op, ltype], dtype=torch.int64), 6, m.LONG, target, _TAG_REQ,
           win.server_comm)


def Win_lock(lock_type: int, rank: int, assertion: int, win) -> None:
    """Begin a passive-target access epoch at ``rank`` (``LOCK_EXCLUSIVE`` or ``LOCK_SHARED``); returns once granted."""
    m = _m()
    if getattr(win, "server", None) is None:
        raise _err(ERR_RMA_SYNC, "Win_lock: the window has no service thread (created before Init?)")
    if lock_type not in (LOCK_EXCLUSIVE, LOCK_SHARED):
        raise _err(13, "Win_lock: lock type")
    if rank in win.locked:
        raise _err(ERR_RMA_SYNC, f"Win_lock: rank {rank} is already locked by this process")
    _request(win, rank, _K_LOCK, ltype=lock_type)
    m.Recv(None, 0, m.BYTE, rank, _TAG_REPLY, win.server_comm)
    win.locked[rank] = lock_type


def Win_unlock(rank: int, win) -> None:
    """End the epoch: returns when the target has applied every operation issued since ``Win_lock``."""
    m = _m()
    if rank not in getattr(win, "locked", {}):
        raise _err(ERR_RMA_SYNC, f"Win_unlock: rank {rank} is not locked")
    _request(win, rank, _K_UNLOCK)
    m.Recv(None, 0, m.BYTE, rank, _TAG_REPLY, win.server_comm)
    del win.locked[rank]


def passive_op(win, kind: str, target: int, disp: int, count: int, dtype: int, data, op) -> bool:
    """Issue ``Put / Accumulate / Get`` at once when ``target`` is locked by this process; False = not in a passive
    epoch (the caller queues the operation for the next fence / complete)."""
    if target not in getattr(win, "locked", {}):
        return False
    m = _m()
    if kind == "get":
        _request(win, target, _K_GET, disp, count, dtype)
        m.Recv(data[:count], count, dtype, target, _TAG_REPLY, win.server_comm)
        return True
    _request(win, target, _K_PUT if kind == "put" else _K_ACC, disp, count, dtype, -1 if op is None else op)
    m.Send(data, count, dtype, target, _TAG_PAY, win.server_comm)
    return True


__all__ = [_n for _n in dir() if _n[0].isupper() and _n not in ("Any", "Callable", "Dict", "List", "Optional", "Sequence",
                                                                 "Tuple")]
