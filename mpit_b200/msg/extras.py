"""The wider MPI surface of the reference's binding, composed from the native primitives.

/root/reference/mpifuncs.c wraps 258 MPI functions; the tree calls 17 of them.  The core of
``mpit_b200.msg`` implements those natively (C++).  This module adds the rest of the chapters a user of the
binding could reasonably have reached for -- send modes, persistent requests, Test/Wait-some/any, the vector
collectives, Exscan/Reduce_scatter, groups and communicator construction, user reduction ops, error classes,
environment queries, and fence-synchronised one-sided windows -- as thin Python compositions of the native
point-to-point and collective calls.  The other chapters live beside it: ``datatypes.py`` (derived datatypes),
``fileio.py`` (MPI-IO), ``topo.py``, ``attrs.py`` (caching, error handlers, Info), ``rma.py`` (post/start/complete/wait)
and ``dynproc.py`` (inter-communicators, ports).

Everything here follows the conventions of the core module: results are returned instead of written through
out-parameters, errors raise :class:`MPIError`.
"""
from __future__ import annotations

import socket
from typing import Callable, Dict, List, Sequence, Tuple

import numpy as np
import torch

# ---- constants (/root/reference/lua-mpi.h:125-231) -------------------------------------------------
PROC_NULL, ROOT = -2, -3
MAX_PROCESSOR_NAME, MAX_ERROR_STRING, MAX_OBJECT_NAME = 256, 512, 128
IDENT, CONGRUENT, SIMILAR, UNEQUAL = 0, 1, 2, 3
(ERR_BUFFER, ERR_COUNT, ERR_TYPE, ERR_TAG, ERR_COMM, ERR_RANK, ERR_REQUEST, ERR_ROOT, ERR_GROUP, ERR_OP, ERR_TOPOLOGY,
 ERR_DIMS, ERR_ARG, ERR_UNKNOWN) = range(1, 15)
ERR_OTHER, ERR_INTERN, ERR_IN_STATUS, ERR_PENDING, ERR_WIN, ERR_RMA_SYNC, ERR_LASTCODE = 16, 17, 18, 19, 45, 50, 92
LXOR, BXOR = 8, 9                                 # the native layer implements SUM..BOR (0..7)
MAXLOC, MINLOC = 10, 11                           # over the {value, int} pair types (FLOAT_INT, 2INT ...)
OP_NULL, GROUP_NULL, DATATYPE_NULL, ERRHANDLER_NULL, INFO_NULL, WIN_NULL = -1, None, -1, -1, -1, None
_ERR_NAMES = {0: "MPI_SUCCESS: no errors", 1: "MPI_ERR_BUFFER: invalid buffer pointer", 2: "MPI_ERR_COUNT: invalid count",
              3: "MPI_ERR_TYPE: invalid datatype", 4: "MPI_ERR_TAG: invalid tag", 5: "MPI_ERR_COMM: invalid communicator",
              6: "MPI_ERR_RANK: invalid rank", 7: "MPI_ERR_REQUEST: invalid request", 8: "MPI_ERR_ROOT: invalid root",
              9: "MPI_ERR_GROUP: invalid group", 10: "MPI_ERR_OP: invalid operation", 13: "MPI_ERR_ARG: invalid argument",
              14: "MPI_ERR_UNKNOWN: unknown error", 15: "MPI_ERR_TRUNCATE: message truncated on receive",
              16: "MPI_ERR_OTHER: known error not in this list", 17: "MPI_ERR_INTERN: internal error",
              45: "MPI_ERR_WIN: invalid window", 50: "MPI_ERR_RMA_SYNC: wrong RMA synchronisation",
              11: "MPI_ERR_TOPOLOGY: invalid topology", 12: "MPI_ERR_DIMS: invalid dimension argument",
              20: "MPI_ERR_ACCESS: permission denied", 21: "MPI_ERR_AMODE: invalid access mode",
              22: "MPI_ERR_BAD_FILE: invalid file name", 23: "MPI_ERR_CONVERSION: data conversion failed",
              24: "MPI_ERR_DUP_DATAREP: data representation already registered", 25: "MPI_ERR_FILE_EXISTS: file exists",
              26: "MPI_ERR_FILE_IN_USE: file is open", 27: "MPI_ERR_FILE: invalid file handle",
              28: "MPI_ERR_INFO: invalid info object", 29: "MPI_ERR_INFO_KEY: invalid info key",
              30: "MPI_ERR_INFO_VALUE: invalid info value", 31: "MPI_ERR_INFO_NOKEY: no such info key",
              32: "MPI_ERR_IO: input/output error", 33: "MPI_ERR_NAME: service name not published",
              35: "MPI_ERR_NOT_SAME: collective arguments differ", 36: "MPI_ERR_NO_SPACE: not enough space",
              37: "MPI_ERR_NO_SUCH_FILE: file does not exist", 38: "MPI_ERR_PORT: invalid port name",
              39: "MPI_ERR_QUOTA: quota exceeded", 40: "MPI_ERR_READ_ONLY: read-only file",
              41: "MPI_ERR_SERVICE: invalid service name", 42: "MPI_ERR_SPAWN: processes cannot be spawned",
              43: "MPI_ERR_UNSUPPORTED_DATAREP: unsupported data representation",
              44: "MPI_ERR_UNSUPPORTED_OPERATION: unsupported operation", 48: "MPI_ERR_KEYVAL: invalid keyval"}
_COLL_TAG = 32100          # tag space of the composed collectives (user tags in the tree are 1..8)


def _m():
    from .. import msg
    return msg


def _flat(buf):
    """1-D view of a user buffer (torch tensor or numpy array) that aliases its memory."""
    if isinstance(buf, torch.Tensor):
        return buf.view(-1) if buf.is_contiguous() else buf.reshape(-1)
    return np.asarray(buf).reshape(-1)


# ---- environment -----------------------------------------------------------------------------------
def Get_version() -> Tuple[int, int]:
    """MPI_Get_version: the standard level whose semantics this layer follows for what it implements."""
    return 3, 1


def Get_library_version() -> str:
    return "mpit_b200 host message layer (TCP, progress thread) -- see csrc/msg/hostcomm.cpp"


def Finalized() -> bool:
    return not _m().Initialized() and getattr(_m(), "_was_finalized", False)


def Query_thread() -> int:
    return _m().THREAD_MULTIPLE       # the native layer is thread-safe (one lock around the matching queues)


def Is_thread_main() -> bool:
    import threading
    return threading.current_thread() is threading.main_thread()


def Error_class(code: int) -> int:
    """Codes of this layer are their own class; a code from ``Add_error_code`` maps to the class it was added to."""
    from . import attrs
    user = attrs.user_error_class(code)
    return code if user is None else user


def Error_string(code: int) -> str:
    from . import attrs
    user = attrs.user_error_string(code)
    return user if user is not None else _ERR_NAMES.get(code, f"error class {code}")


def Type_size(dtype: int) -> int:
    from . import datatypes
    return datatypes.Type_size(dtype)


def Type_get_extent(dtype: int) -> Tuple[int, int]:
    """-> (lower bound, extent), basic and derived datatypes alike."""
    from . import datatypes
    return datatypes.Type_get_extent(dtype)


def Alloc_mem(nbytes: int) -> torch.Tensor:
    """MPI_Alloc_mem: page-locked host memory when CUDA is present (fast staging for device tensors)."""
    t = torch.empty(nbytes, dtype=torch.uint8)
    return t.pin_memory() if torch.cuda.is_available() else t


def Free_mem(buf) -> None:
    del buf


def Get_hostname() -> str:
    return socket.gethostname()


# ---- send modes (mpifuncs.c Bsend/Rsend/Ssend and their I-variants) ---------------------------------
# The native Send returns once the message is handed to the progress thread (buffered semantics), which is a
# legal implementation of every mode except the synchronous one; Ssend adds the handshake.
def Bsend(buf, count, dtype, dest, tag, comm=0):
    _m().Send(buf, count, dtype, dest, tag, comm)


Rsend = Bsend


def Ssend(buf, count, dtype, dest, tag, comm=0):
    """Synchronous send: returns only after the receiver has matched the message (0-byte acknowledgement)."""
    m = _m()
    m.Send(buf, count, dtype, dest, tag, comm)
    m.Recv(None, 0, m.BYTE, dest, _COLL_TAG + 90, comm)


def Ssend_ack(source, comm=0):
    """Receiver side of :func:`Ssend` (call after the matching Recv)."""
    m = _m()
    m.Send(None, 0, m.BYTE, source, _COLL_TAG + 90, comm)


def Ibsend(buf, count, dtype, dest, tag, comm, request, i=0):
    _m().Isend(buf, count, dtype, dest, tag, comm, request, i)


Irsend = Issend = Ibsend


# ---- completion of several requests ----------------------------------------------------------------
def Testany(request, status=None) -> Tuple[int, bool]:
    """-> (index, flag).  index is UNDEFINED when no active request exists or none has completed."""
    m = _m()
    active = [i for i in range(len(request)) if not request.is_null(i)]
    if not active:
        return m.UNDEFINED, True
    for i in active:
        if m.Test(request, status, i):
            return i, True
    return m.UNDEFINED, False


def Testsome(request, statuses=None) -> List[int]:
    """Indices that completed now (empty list: none yet; UNDEFINED semantics -> ``None`` when nothing is active)."""
    m = _m()
    active = [i for i in range(len(request)) if not request.is_null(i)]
    if not active:
        return None
    return [i for i in active if m.Test(request, None, i)]


def Waitsome(request, statuses=None) -> List[int]:
    import time
    while True:
        done = Testsome(request, statuses)
        if done is None or done:
            return done or []
        time.sleep(0)


# ---- persistent requests (Send_init / Recv_init / Start / Startall / Request_free) ------------------
def _persist(request) -> Dict[int, tuple]:
    if not hasattr(request, "_persistent"):
        request._persistent = {}
    return request._persistent


def Send_init(buf, count, dtype, dest, tag, comm, request, i=0):
    _persist(request)[i] = ("send", buf, count, dtype, dest, tag, comm)


Bsend_init = Rsend_init = Ssend_init = Send_init


def Recv_init(buf, count, dtype, source, tag, comm, request, i=0):
    _persist(request)[i] = ("recv", buf, count, dtype, source, tag, comm)


def Start(request, i=0):
    m = _m()
    kind, buf, count, dtype, peer, tag, comm = _persist(request)[i]
    (m.Isend if kind == "send" else m.Irecv)(buf, count, dtype, peer, tag, comm, request, i)


def Startall(request):
    for i in sorted(_persist(request)):
        Start(request, i)


def Request_free(request, i=0):
    _persist(request).pop(i, None)
    request._arr[i] = _m().REQUEST_NULL


# ---- vector collectives ----------------------------------------------------------------------------
def _p2p_exchange(sends, recvs, comm):
    """sends/recvs: lists of (peer, flat_buffer, count, dtype).  All posted non-blocking, then waited."""
    m = _m()
    req = m.Request(len(sends) + len(recvs))
    k = 0
    for peer, b, cnt, dt in recvs:
        m.Irecv(b, cnt, dt, peer, _COLL_TAG, comm, req, k)
        k += 1
    for peer, b, cnt, dt in sends:
        m.Isend(b, cnt, dt, peer, _COLL_TAG, comm, req, k)
        k += 1
    m.Waitall(req)


def Gatherv(sbuf, scount, dtype, rbuf, rcounts: Sequence[int], displs: Sequence[int], root, comm=0):
    m = _m()
    me, n = m.Comm_rank(comm), m.Comm_size(comm)
    s = _flat(sbuf)
    if me != root:
        m.Send(s[:scount], scount, dtype, root, _COLL_TAG, comm)
        return
    r = _flat(rbuf)
    recvs = [(p, r[displs[p]:displs[p] + rcounts[p]], rcounts[p], dtype) for p in range(n) if p != root]
    r[displs[root]:displs[root] + rcounts[root]] = s[:rcounts[root]]
    _p2p_exchange([], recvs, comm)


def Scatterv(sbuf, scounts: Sequence[int], displs: Sequence[int], dtype, rbuf, rcount, root, comm=0):
    m = _m()
    me, n = m.Comm_rank(comm), m.Comm_size(comm)
    r = _flat(rbuf)
    if me != root:
        m.Recv(r[:rcount], rcount, dtype, root, _COLL_TAG, comm)
        return
    s = _flat(sbuf)
    sends = [(p, s[displs[p]:displs[p] + scounts[p]], scounts[p], dtype) for p in range(n) if p != root]
    r[:scounts[root]] = s[displs[root]:displs[root] + scounts[root]]
    _p2p_exchange(sends, [], comm)


def Allgatherv(sbuf, scount, dtype, rbuf, rcounts: Sequence[int], displs: Sequence[int], comm=0):
    m = _m()
    me, n = m.Comm_rank(comm), m.Comm_size(comm)
    s, r = _flat(sbuf), _flat(rbuf)
    r[displs[me]:displs[me] + rcounts[me]] = s[:rcounts[me]]
    sends = [(p, s[:scount], scount, dtype) for p in range(n) if p != me]
    recvs = [(p, r[displs[p]:displs[p] + rcounts[p]], rcounts[p], dtype) for p in range(n) if p != me]
    _p2p_exchange(sends, recvs, comm)


def Alltoallv(sbuf, scounts, sdispls, dtype, rbuf, rcounts, rdispls, comm=0):
    m = _m()
    me, n = m.Comm_rank(comm), m.Comm_size(comm)
    s, r = _flat(sbuf), _flat(rbuf)
    r[rdispls[me]:rdispls[me] + rcounts[me]] = s[sdispls[me]:sdispls[me] + scounts[me]]
    sends = [(p, s[sdispls[p]:sdispls[p] + scounts[p]], scounts[p], dtype) for p in range(n) if p != me]
    recvs = [(p, r[rdispls[p]:rdispls[p] + rcounts[p]], rcounts[p], dtype) for p in range(n) if p != me]
    _p2p_exchange(sends, recvs, comm)


def Exscan(sbuf, rbuf, count, dtype, op, comm=0):
    """Exclusive prefix reduction: rank r gets op(s_0..s_{r-1}); rbuf on rank 0 is left untouched."""
    m = _m()
    me, n = m.Comm_rank(comm), m.Comm_size(comm)
    s = _flat(sbuf)
    tmp = s.clone() if isinstance(s, torch.Tensor) else s.copy()
    m.Scan(sbuf, tmp, count, dtype, op, comm)          # inclusive; shift it one rank to the right
    r = _flat(rbuf)
    sends = [(me + 1, tmp[:count], count, dtype)] if me + 1 < n else []
    recvs = [(me - 1, r[:count], count, dtype)] if me > 0 else []
    _p2p_exchange(sends, recvs, comm)


def Reduce_scatter(sbuf, rbuf, rcounts: Sequence[int], dtype, op, comm=0):
    """General (uneven) reduce-scatter: reduce to rank 0, scatter the pieces."""
    m = _m()
    me = m.Comm_rank(comm)
    total = int(sum(rcounts))
    s = _flat(sbuf)
    full = (torch.empty_like(s[:total]) if isinstance(s, torch.Tensor) else np.empty_like(s[:total]))
    m.Reduce(s[:total], full, total, dtype, op, 0, comm)
    displs = [int(sum(rcounts[:i])) for i in range(len(rcounts))]
    Scatterv(full, list(rcounts), displs, dtype, rbuf, rcounts[me], 0, comm)


# ---- user reduction operations (Op_create / Op_free) ------------------------------------------------
_USER_OPS: Dict[int, Tuple[Callable, bool]] = {}
_USER_OP_BASE = 100


def Op_create(fn: Callable, commute: bool = True) -> int:
    """``fn(invec, inoutvec)`` combines in place into ``inoutvec`` (the MPI_User_function contract)."""
    h = _USER_OP_BASE + len(_USER_OPS)
    _USER_OPS[h] = (fn, commute)
    return h


def Op_free(op: int) -> None:
    if op >= _USER_OP_BASE:
        _USER_OPS.pop(op, None)


def _lxor(invec, inout):
    inout[:] = ((invec != 0) ^ (inout != 0)).to(inout.dtype) if isinstance(inout, torch.Tensor) else \
        ((invec != 0) ^ (inout != 0)).astype(inout.dtype)


def _bxor(invec, inout):
    inout ^= invec


_USER_OPS[LXOR] = (_lxor, True)      # the two predefined ops the native reducer does not implement
_USER_OPS[BXOR] = (_bxor, True)


def is_user_op(op: int) -> bool:
    return op in _USER_OPS


def reduce_user_op(sbuf, rbuf, count, dtype, op, root, comm, everyone: bool):
    """Reduce/Allreduce with an Op_create'd operation: gather to the root, fold in rank order, (broadcast)."""
    m = _m()
    me, n = m.Comm_rank(comm), m.Comm_size(comm)
    fn, _ = _USER_OPS[op]
    s = _flat(sbuf)
    root_ = 0 if everyone else root
    gathered = None
    if me == root_:
        gathered = torch.empty(n * count, dtype=s.dtype) if isinstance(s, torch.Tensor) else np.empty(n * count, s.dtype)
    m.Gather(s[:count], count, dtype, gathered, root_, comm)
    r = _flat(rbuf) if rbuf is not None else None
    if me == root_:
        acc = gathered[:count].clone() if isinstance(gathered, torch.Tensor) else gathered[:count].copy()
        for p in range(1, n):
            # MPI: inoutvec = invec op inoutvec with invec the lower-ranked operand
            nxt = gathered[p * count:(p + 1) * count]
            nxt = nxt.clone() if isinstance(nxt, torch.Tensor) else nxt.copy()
            fn(acc, nxt)
            acc = nxt
        r[:count] = acc
    if everyone:
        m.Bcast(r[:count], count, dtype, root_, comm)


def needs_packed_reduce(dtype: int, op: int) -> bool:
    from . import datatypes
    return op in (MAXLOC, MINLOC) or dtype >= datatypes.DERIVED_BASE


def _fold(a: np.ndarray, b: np.ndarray, op: int) -> np.ndarray:
    """``a op b`` with ``a`` the lower-ranked operand; record arrays are the pair types of MAXLOC / MINLOC."""
    m = _m()
    if op in (MAXLOC, MINLOC):
        if a.dtype.names is None or len(a.dtype.names) != 2:
            raise m.MPIError(ERR_OP, "MAXLOC / MINLOC need a {value, int} pair datatype")
        v, i = a.dtype.names
        better = (a[v] > b[v]) if op == MAXLOC else (a[v] < b[v])
        tie = a[v] == b[v]
        out = b.copy()
        take_a = better | (tie & (a[i] < b[i]))
        out[take_a] = a[take_a]
        return out
    if a.dtype.names is not None:
        raise m.MPIError(ERR_OP, "arithmetic reduction of a record datatype")
    table = {m.SUM: np.add, m.PROD: np.multiply, m.MAX: np.maximum, m.MIN: np.minimum, m.BAND: np.bitwise_and,
             m.BOR: np.bitwise_or, BXOR: np.bitwise_xor}
    if op in table:
        return table[op](a, b)
    if op in (m.LAND, m.LOR, LXOR):
        f = {m.LAND: np.logical_and, m.LOR: np.logical_or, LXOR: np.logical_xor}[op]
        return f(a != 0, b != 0).astype(a.dtype)
    raise m.MPIError(ERR_OP, f"reduction op {op}")


def reduce_packed(sbuf, rbuf, count, dtype, op, root, comm, everyone: bool):
    """Reduce / Allreduce for datatypes or operations the native reducer has no code for (UNSIGNED_SHORT,
    UNSIGNED_LONG, LONG_DOUBLE, the pair types with MAXLOC / MINLOC, contiguous derived types of one element kind):
    the packed elements are gathered on the root, folded in rank order with NumPy, and broadcast when everyone
    wants the result."""
    from . import datatypes
    m = _m()
    me, n = m.Comm_rank(comm), m.Comm_size(comm)
    if is_user_op(op) and op not in (LXOR, BXOR):
        raise m.MPIError(ERR_OP, "user-defined operations take basic datatypes")
    el = datatypes.packed_numpy_dtype(dtype, records=op in (MAXLOC, MINLOC))   # MAXLOC sees (value, index) records
    # argument errors must surface on every rank BEFORE anyone communicates
    if op in (MAXLOC, MINLOC):
        if el.names is None or len(el.names) != 2:
            raise m.MPIError(ERR_OP, "MAXLOC / MINLOC need a {value, int} pair datatype")
    elif el.names is not None:
        raise m.MPIError(ERR_OP, "arithmetic reduction of a record datatype")
    elif op not in (m.SUM, m.PROD, m.MAX, m.MIN, m.BAND, m.BOR, BXOR, m.LAND, m.LOR, LXOR):
        raise m.MPIError(ERR_OP, f"reduction op {op}")
    mine = datatypes.pack_bytes(sbuf, count, dtype)
    root_ = 0 if everyone else root
    gathered = np.empty(n * mine.size, dtype=np.uint8) if me == root_ else None
    m.Gather(mine, mine.size, m.BYTE, gathered, root_, comm)
    result = np.empty(mine.size, dtype=np.uint8)
    if me == root_:
        parts = gathered.reshape(n, mine.size)
        acc = parts[0].view(el).copy()
        for p in range(1, n):
            acc = _fold(acc, parts[p].view(el), op)
        result[:] = acc.view(np.uint8)
    if everyone:
        m.Bcast(result, result.size, m.BYTE, root_, comm)
    if everyone or me == root_:
        datatypes.unpack_bytes(result, rbuf, count, dtype)


# ---- groups and communicator construction -----------------------------------------------------------
class Group:
    """An ordered set of world ranks (MPI_Group)."""

    def __init__(self, ranks: Sequence[int]):
        self.ranks = list(ranks)

    def __eq__(self, other):
        return isinstance(other, Group) and self.ranks == other.ranks


GROUP_EMPTY = Group([])


def Comm_group(comm=0) -> Group:
    m = _m()
    n = m.Comm_size(comm)
    mine = torch.tensor([m.Comm_rank(m.COMM_WORLD)], dtype=torch.int32)
    allr = torch.empty(n, dtype=torch.int32)
    m.Allgather(mine, 1, m.INT, allr, comm)
    return Group(allr.tolist())


def Group_size(g: Group) -> int:
    return len(g.ranks)


def Group_rank(g: Group) -> int:
    me = _m().Comm_rank(_m().COMM_WORLD)
    return g.ranks.index(me) if me in g.ranks else _m().UNDEFINED


def Group_incl(g: Group, ranks: Sequence[int]) -> Group:
    return Group([g.ranks[i] for i in ranks])


def Group_excl(g: Group, ranks: Sequence[int]) -> Group:
    drop = set(ranks)
    return Group([r for i, r in enumerate(g.ranks) if i not in drop])


def Group_union(a: Group, b: Group) -> Group:
    return Group(a.ranks + [r for r in b.ranks if r not in a.ranks])


def Group_intersection(a: Group, b: Group) -> Group:
    return Group([r for r in a.ranks if r in b.ranks])


def Group_difference(a: Group, b: Group) -> Group:
    return Group([r for r in a.ranks if r not in b.ranks])


def Group_translate_ranks(a: Group, ranks: Sequence[int], b: Group) -> List[int]:
    und = _m().UNDEFINED
    return [b.ranks.index(a.ranks[i]) if a.ranks[i] in b.ranks else und for i in ranks]


def Group_compare(a: Group, b: Group) -> int:
    if a.ranks == b.ranks:
        return IDENT
    return SIMILAR if sorted(a.ranks) == sorted(b.ranks) else UNEQUAL


def Group_free(g: Group) -> None:
    g.ranks = []


def Comm_create(comm, group: Group) -> int:
    """Collective over ``comm``: the new communicator holds ``group`` in its order; others get COMM_NULL."""
    m = _m()
    me = m.Comm_rank(m.COMM_WORLD)
    if me in group.ranks:
        return m.Comm_split(comm, 0, group.ranks.index(me))
    return m.Comm_split(comm, m.UNDEFINED, 0)


def Comm_compare(c1, c2) -> int:
    if c1 == c2:
        return IDENT
    g = Group_compare(Comm_group(c1), Comm_group(c2))
    return CONGRUENT if g == IDENT else g


def Comm_test_inter(comm) -> bool:
    from . import dynproc
    return isinstance(comm, dynproc.Intercomm)


_COMM_NAMES: Dict[int, str] = {0: "MPI_COMM_WORLD", 1: "MPI_COMM_SELF"}


def Comm_set_name(comm, name: str) -> None:
    _COMM_NAMES[comm] = name[:MAX_OBJECT_NAME]


def Comm_get_name(comm) -> str:
    return _COMM_NAMES.get(comm, "")


# ---- one-sided communication with fence synchronisation (host memory) -------------------------------
class Win:
    """``MPI_Win_create`` over a host tensor.  Put / Get / Accumulate are queued at the origin and performed at the
    next :func:`Win_fence`, which is when MPI guarantees them complete under active-target synchronisation
    (/root/reference/mpifuncs.c:2337,1656,1131,9,2367).  Inside a passive-target epoch (``Win_lock`` .. ``Win_unlock``,
    rma.py) they are sent at once to the target's service thread instead.  The device-memory counterpart is
    :class:`mpit_b200.parallel.DeviceWindow`."""

    def __init__(self, base, disp_unit: int, comm: int):
        self.base = _flat(base)
        self.disp_unit, self.comm = disp_unit, comm
        self.ops: List[tuple] = []
        # Win_create is collective, so the communicator's group can be gathered here once; the post / start /
        # complete / wait epochs of rma.py need it and are not collective themselves
        self.group = Comm_group(comm) if _m().Initialized() else Group([0])


def Win_create(base, size_bytes: int, disp_unit: int, info=None, comm=0) -> Win:
    """Collective.  Also starts the window's passive-target service thread (rma.py)."""
    from . import rma
    win = Win(base, disp_unit, comm)
    rma.start_passive_server(win)
    return win


def Put(origin, count, dtype, target_rank, target_disp, win: Win):
    from . import rma
    o = _flat(origin)[:count]
    if rma.passive_op(win, "put", target_rank, target_disp, count, dtype, o, None):
        return
    win.ops.append(("put", target_rank, target_disp, count, dtype, o.clone() if isinstance(o, torch.Tensor) else o.copy(), None))


def Accumulate(origin, count, dtype, target_rank, target_disp, op, win: Win):
    from . import rma
    o = _flat(origin)[:count]
    if rma.passive_op(win, "acc", target_rank, target_disp, count, dtype, o, op):
        return
    win.ops.append(("acc", target_rank, target_disp, count, dtype, o.clone() if isinstance(o, torch.Tensor) else o.copy(), op))


def Get(origin, count, dtype, target_rank, target_disp, win: Win):
    from . import rma
    if rma.passive_op(win, "get", target_rank, target_disp, count, dtype, _flat(origin), None):
        return
    win.ops.append(("get", target_rank, target_disp, count, dtype, _flat(origin), None))


def _apply_op(dst, src, op):
    m = _m()
    if op == m.SUM:
        dst += src
    elif op == m.PROD:
        dst *= src
    elif op == m.MAX:
        dst[:] = torch.maximum(dst, src) if isinstance(dst, torch.Tensor) else np.maximum(dst, src)
    elif op == m.MIN:
        dst[:] = torch.minimum(dst, src) if isinstance(dst, torch.Tensor) else np.minimum(dst, src)
    elif is_user_op(op):
        fn, _ = _USER_OPS[op]
        tmp = dst.clone() if isinstance(dst, torch.Tensor) else dst.copy()
        fn(src, tmp)
        dst[:] = tmp
    else:
        raise _m().MPIError(ERR_OP, "Accumulate")


def Win_fence(assertion: int, win: Win) -> None:
    """Close the access/exposure epoch: exchange the queued operations and apply them at their targets."""
    m = _m()
    comm = win.comm
    me, n = m.Comm_rank(comm), m.Comm_size(comm)
    # 1. everybody learns how many descriptors it will receive from everybody
    out_cnt = torch.zeros(n, dtype=torch.int64)
    for op in win.ops:
        out_cnt[op[1]] += 1
    in_cnt = torch.zeros(n, dtype=torch.int64)
    m.Alltoall(out_cnt, 1, m.LONG, in_cnt, comm)
    # 2. descriptors {kind, disp, count, dtype, op} then payloads, pairwise and in issue order
    kinds = {"put": 0, "acc": 1, "get": 2}
    local = [op for op in win.ops if op[1] == me]
    req = m.Request(1)
    for peer in range(n):
        if peer == me:
            continue
        mine = [op for op in win.ops if op[1] == peer]
        desc = torch.tensor([[kinds[o[0]], o[2], o[3], o[4], -1 if o[6] is None else o[6]] for o in mine],
                            dtype=torch.int64).reshape(-1)
        theirs = torch.empty(int(in_cnt[peer]) * 5, dtype=torch.int64)
        m.Irecv(theirs, theirs.numel(), m.LONG, peer, _COLL_TAG + 1, comm, req)
        m.Send(desc, desc.numel(), m.LONG, peer, _COLL_TAG + 1, comm)
        m.Wait(req)
        def serve_theirs():
            t = theirs.view(-1, 5).tolist()
            for kind, disp, cnt, dt, op in t:
                if kind == 2:
                    continue
                tmp = torch.empty(cnt, dtype=win.base.dtype) if isinstance(win.base, torch.Tensor) else np.empty(cnt, win.base.dtype)
                m.Recv(tmp, cnt, dt, peer, _COLL_TAG + 2, comm)
                if kind == 0:
                    win.base[disp:disp + cnt] = tmp
                else:
                    _apply_op(win.base[disp:disp + cnt], tmp, op)
            for kind, disp, cnt, dt, op in t:
                if kind == 2:
                    m.Send(win.base[disp:disp + cnt], cnt, dt, peer, _COLL_TAG + 3, comm)

        # puts/accumulates travel eagerly (the native Send returns once the progress thread owns the message), so
        # both sides can send first; gets are served after the peer's puts, which makes a Get that follows a Put to
        # the same location inside one epoch well defined here
        for o in mine:
            if o[0] != "get":
                m.Send(o[5], o[3], o[4], peer, _COLL_TAG + 2, comm)
        serve_theirs()
        for o in mine:
            if o[0] == "get":
                m.Recv(o[5][:o[3]], o[3], o[4], peer, _COLL_TAG + 3, comm)
    # 3. operations that target myself
    for kind, _, disp, cnt, dt, data, op in local:
        if kind == "put":
            win.base[disp:disp + cnt] = data
        elif kind == "acc":
            _apply_op(win.base[disp:disp + cnt], data, op)
        else:
            data[:cnt] = win.base[disp:disp + cnt]
    win.ops = []
    m.Barrier(comm)


def Win_free(win: Win) -> None:
    """Collective: stops the service thread, drops cached attributes."""
    from . import attrs, rma
    rma.stop_passive_server(win)
    attrs.drop_object("win", win)
    win.ops = []
    win.base = None
