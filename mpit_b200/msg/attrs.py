"""Caching (keyvals / attributes), error handlers, the Info object and the small "external interfaces" calls.

These are the chapters of /root/reference/mpifuncs.c that are pure bookkeeping around a handle (SURVEY.md 2.1, N7:
"Comm/Group/attr 44/46", "Info 9/9", "Env 25/29", "Ext 6/7"): ``Comm/Type/Win_create_keyval``, ``*_set/get/delete_
attr``, ``*_create/set/get/call_errhandler``, ``Add_error_class/code/string``, ``Info_*``, generalized requests,
``Status_set_*``, ``Buffer_attach/detach``, ``Op_commutative``, ``Reduce_local``, ``Pcontrol``, ``Init_thread``,
``Request_get_status``, ``Sendrecv_replace``, ``Alltoallw``, ``Group_range_incl/excl``.  None of them is called by a
script of the reference.  They hold no MPI-library state here either: tables keyed by the handle.

Values MPI returns through out-parameters are returned; "flag" results come back as ``(value, flag)``.
"""
from __future__ import annotations

import ctypes
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

ERR_ARG, ERR_KEYVAL, ERR_INFO_KEY, ERR_INFO_NOKEY, ERR_INFO_VALUE, ERR_INFO, ERR_OP, ERR_BUFFER = 13, 48, 29, 31, 30, 28, 10, 1
MAX_INFO_KEY, MAX_INFO_VAL = 255, 1024
KEYVAL_INVALID = -1
# predefined attribute keys
TAG_UB, HOST, IO, WTIME_IS_GLOBAL, UNIVERSE_SIZE, APPNUM, LASTUSEDCODE = 1, 2, 3, 4, 5, 6, 7
WIN_BASE, WIN_SIZE, WIN_DISP_UNIT = 8, 9, 10
BSEND_OVERHEAD = 96
ERRORS_ARE_FATAL, ERRORS_RETURN = 0, 1
ERR_LASTCODE_BASE = 92


def _m():
    from .. import msg
    return msg


def _err(code: int, what: str):
    return _m().MPIError(code, what)


# ---- keyvals and attributes ----------------------------------------------------------------------------------------
# one table for the three object kinds; a keyval remembers which kind it was created for
_KEYVALS: Dict[int, Tuple[str, Optional[Callable], Optional[Callable], Any]] = {}
_ATTRS: Dict[Tuple[str, Any], Dict[int, Any]] = {}
_next_keyval = [100]


def _obj_key(kind: str, obj) -> Tuple[str, Any]:
    return kind, (obj if isinstance(obj, int) else id(obj))


def _create_keyval(kind: str, copy_fn, delete_fn, extra_state) -> int:
    k = _next_keyval[0]
    _next_keyval[0] += 1
    _KEYVALS[k] = (kind, copy_fn, delete_fn, extra_state)
    return k


def _free_keyval(kind: str, keyval: int) -> int:
    if keyval not in _KEYVALS or _KEYVALS[keyval][0] != kind:
        raise _err(ERR_KEYVAL, f"{kind} keyval {keyval}")
    del _KEYVALS[keyval]           # attributes already cached under it stay readable until deleted, as in MPI
    return KEYVAL_INVALID


def _predefined(kind: str, obj, keyval: int):
    m = _m()
    if kind == "comm":
        table = {TAG_UB: 2 ** 30 - 1, HOST: m.PROC_NULL, IO: m.ANY_SOURCE, WTIME_IS_GLOBAL: 0,
                 UNIVERSE_SIZE: m.Comm_size(m.COMM_WORLD) if m.Initialized() else 1, APPNUM: 0,
                 LASTUSEDCODE: ERR_LASTCODE_BASE + len(_USER_ERR_STRINGS)}
        return table.get(keyval)
    if kind == "win" and keyval in (WIN_BASE, WIN_SIZE, WIN_DISP_UNIT):
        base = getattr(obj, "base", None)
        if keyval == WIN_BASE:
            return base
        if keyval == WIN_SIZE:
            return 0 if base is None else base.numel() * base.element_size() if isinstance(base, torch.Tensor) \
                else base.nbytes
        return getattr(obj, "disp_unit", 1)
    return None


def _set_attr(kind: str, obj, keyval: int, value) -> None:
    if keyval not in _KEYVALS or _KEYVALS[keyval][0] != kind:
        raise _err(ERR_KEYVAL, f"{kind} keyval {keyval}")
    d = _ATTRS.setdefault(_obj_key(kind, obj), {})
    if keyval in d:
        _delete_attr(kind, obj, keyval)
        d = _ATTRS.setdefault(_obj_key(kind, obj), {})
    d[keyval] = value


def _get_attr(kind: str, obj, keyval: int) -> Tuple[Any, bool]:
    pre = _predefined(kind, obj, keyval)
    if pre is not None:
        return pre, True
    d = _ATTRS.get(_obj_key(kind, obj), {})
    return (d[keyval], True) if keyval in d else (None, False)


def _delete_attr(kind: str, obj, keyval: int) -> None:
    d = _ATTRS.get(_obj_key(kind, obj), {})
    if keyval not in d:
        raise _err(ERR_KEYVAL, f"no attribute under keyval {keyval}")
    value = d.pop(keyval)
    kv = _KEYVALS.get(keyval)
    if kv and kv[2]:
        kv[2](obj, keyval, value, kv[3])


def copy_attrs(kind: str, old, new) -> None:
    """Attribute propagation of ``Comm_dup`` / ``Type_dup``: the keyval's copy callback decides.
    ``copy_fn(oldobj, keyval, extra_state, value) -> (new_value, flag)``; no callback = not copied (NULL_COPY_FN)."""
    for keyval, value in list(_ATTRS.get(_obj_key(kind, old), {}).items()):
        kv = _KEYVALS.get(keyval)
        if not kv or kv[1] is None:
            continue
        new_value, flag = kv[1](old, keyval, kv[3], value)
        if flag:
            _ATTRS.setdefault(_obj_key(kind, new), {})[keyval] = new_value


def drop_object(kind: str, obj) -> None:
    """Object destruction: run the delete callbacks of everything cached on it."""
    for keyval in list(_ATTRS.get(_obj_key(kind, obj), {})):
        _delete_attr(kind, obj, keyval)
    _ATTRS.pop(_obj_key(kind, obj), None)


def DUP_FN(old, keyval, extra_state, value):
    """MPI_COMM_DUP_FN / MPI_TYPE_DUP_FN / MPI_WIN_DUP_FN: copy the attribute value as is."""
    return value, True


NULL_COPY_FN = None
NULL_DELETE_FN = None


def Comm_create_keyval(copy_fn=None, delete_fn=None, extra_state=None) -> int:
    return _create_keyval("comm", copy_fn, delete_fn, extra_state)


def Comm_free_keyval(keyval: int) -> int:
    return _free_keyval("comm", keyval)


def Comm_set_attr(comm: int, keyval: int, value) -> None:
    _set_attr("comm", comm, keyval, value)


def Comm_get_attr(comm: int, keyval: int) -> Tuple[Any, bool]:
    return _get_attr("comm", comm, keyval)


def Comm_delete_attr(comm: int, keyval: int) -> None:
    _delete_attr("comm", comm, keyval)


def Type_create_keyval(copy_fn=None, delete_fn=None, extra_state=None) -> int:
    return _create_keyval("type", copy_fn, delete_fn, extra_state)


def Type_free_keyval(keyval: int) -> int:
    return _free_keyval("type", keyval)


def Type_set_attr(dtype: int, keyval: int, value) -> None:
    _set_attr("type", dtype, keyval, value)


def Type_get_attr(dtype: int, keyval: int) -> Tuple[Any, bool]:
    return _get_attr("type", dtype, keyval)


def Type_delete_attr(dtype: int, keyval: int) -> None:
    _delete_attr("type", dtype, keyval)


def Win_create_keyval(copy_fn=None, delete_fn=None, extra_state=None) -> int:
    return _create_keyval("win", copy_fn, delete_fn, extra_state)


def Win_free_keyval(keyval: int) -> int:
    return _free_keyval("win", keyval)


def Win_set_attr(win, keyval: int, value) -> None:
    _set_attr("win", win, keyval, value)


def Win_get_attr(win, keyval: int) -> Tuple[Any, bool]:
    return _get_attr("win", win, keyval)


def Win_delete_attr(win, keyval: int) -> None:
    _delete_attr("win", win, keyval)


def Win_set_name(win, name: str) -> None:
    win.name = name[:128]


def Win_get_name(win) -> str:
    return getattr(win, "name", "")


# ---- error handlers and user error classes -------------------------------------------------------------------------
_ERRHANDLERS: Dict[int, Tuple[str, Callable]] = {}
_OBJ_ERRH: Dict[Tuple[str, Any], int] = {}
_next_errh = [10]
_USER_ERR_CLASSES: List[int] = []
_USER_ERR_CODES: Dict[int, int] = {}          # code -> class
_USER_ERR_STRINGS: Dict[int, str] = {}


def _create_errhandler(kind: str, fn: Callable) -> int:
    h = _next_errh[0]
    _next_errh[0] += 1
    _ERRHANDLERS[h] = (kind, fn)
    return h


def _call_errhandler(kind: str, obj, errorcode: int) -> None:
    """ERRORS_ARE_FATAL aborts the job, ERRORS_RETURN does nothing (the caller sees the code), a user handler is
    called as ``fn(obj, errorcode)``.  Files default to ERRORS_RETURN, everything else to ERRORS_ARE_FATAL."""
    h = _OBJ_ERRH.get(_obj_key(kind, obj), ERRORS_RETURN if kind == "file" else ERRORS_ARE_FATAL)
    if h == ERRORS_RETURN:
        return
    if h == ERRORS_ARE_FATAL:
        _m().Abort(_m().COMM_WORLD, errorcode)
        raise SystemExit(errorcode)
    _ERRHANDLERS[h][1](obj, errorcode)


def _set_errhandler(kind: str, obj, errhandler: int) -> None:
    if errhandler not in (ERRORS_ARE_FATAL, ERRORS_RETURN) and _ERRHANDLERS.get(errhandler, ("",))[0] != kind:
        raise _err(ERR_ARG, f"error handler {errhandler} was not created for a {kind}")
    _OBJ_ERRH[_obj_key(kind, obj)] = errhandler


def _get_errhandler(kind: str, obj) -> int:
    return _OBJ_ERRH.get(_obj_key(kind, obj), ERRORS_RETURN if kind == "file" else ERRORS_ARE_FATAL)


def Comm_create_errhandler(fn: Callable) -> int:
    return _create_errhandler("comm", fn)


def Comm_set_errhandler(comm: int, errhandler: int) -> None:
    _set_errhandler("comm", comm, errhandler)


def Comm_get_errhandler(comm: int) -> int:
    return _get_errhandler("comm", comm)


def Comm_call_errhandler(comm: int, errorcode: int) -> None:
    _call_errhandler("comm", comm, errorcode)


def Win_create_errhandler(fn: Callable) -> int:
    return _create_errhandler("win", fn)


def Win_set_errhandler(win, errhandler: int) -> None:
    _set_errhandler("win", win, errhandler)


def Win_get_errhandler(win) -> int:
    return _get_errhandler("win", win)


def Win_call_errhandler(win, errorcode: int) -> None:
    _call_errhandler("win", win, errorcode)


def File_create_errhandler(fn: Callable) -> int:
    return _create_errhandler("file", fn)


def File_set_errhandler(fh, errhandler: int) -> None:
    """``fh`` may be FILE_NULL (None): that sets the default for files opened afterwards, as in MPI."""
    _set_errhandler("file", 0 if fh is None else fh, errhandler)


def File_get_errhandler(fh) -> int:
    if fh is not None and _obj_key("file", fh) in _OBJ_ERRH:
        return _OBJ_ERRH[_obj_key("file", fh)]
    return _OBJ_ERRH.get(("file", 0), ERRORS_RETURN)


def File_call_errhandler(fh, errorcode: int) -> None:
    h = File_get_errhandler(fh)
    if h == ERRORS_RETURN:
        return
    if h == ERRORS_ARE_FATAL:
        _m().Abort(_m().COMM_WORLD, errorcode)
        raise SystemExit(errorcode)
    _ERRHANDLERS[h][1](fh, errorcode)


def Errhandler_free(errhandler: int) -> int:
    if errhandler not in (ERRORS_ARE_FATAL, ERRORS_RETURN):
        _ERRHANDLERS.pop(errhandler, None)
    return -1                                                    # ERRHANDLER_NULL


def Add_error_class() -> int:
    c = ERR_LASTCODE_BASE + 1 + len(_USER_ERR_CLASSES) + len(_USER_ERR_CODES)
    _USER_ERR_CLASSES.append(c)
    _USER_ERR_STRINGS.setdefault(c, "")
    return c


def Add_error_code(errorclass: int) -> int:
    c = ERR_LASTCODE_BASE + 1 + len(_USER_ERR_CLASSES) + len(_USER_ERR_CODES)
    _USER_ERR_CODES[c] = errorclass
    _USER_ERR_STRINGS.setdefault(c, "")
    return c


def Add_error_string(errorcode: int, string: str) -> None:
    if errorcode <= ERR_LASTCODE_BASE:
        raise _err(ERR_ARG, "Add_error_string: only user-defined codes can be given a string")
    _USER_ERR_STRINGS[errorcode] = string


def user_error_class(code: int) -> Optional[int]:
    if code in _USER_ERR_CODES:
        return _USER_ERR_CODES[code]
    return code if code in _USER_ERR_CLASSES else None


def user_error_string(code: int) -> Optional[str]:
    return _USER_ERR_STRINGS.get(code)


# ---- the Info object -----------------------------------------------------------------------------------------------
class Info:
    """An ordered set of (key, value) string pairs (MPI_Info)."""

    def __init__(self, items: Optional[Dict[str, str]] = None):
        self.items: Dict[str, str] = dict(items or {})


INFO_NULL = None


def Info_create() -> Info:
    return Info()


def Info_set(info: Info, key: str, value: str) -> None:
    if not key or len(key) > MAX_INFO_KEY:
        raise _err(ERR_INFO_KEY, "Info_set")
    if len(value) > MAX_INFO_VAL:
        raise _err(ERR_INFO_VALUE, "Info_set")
    info.items[key] = value


def Info_get(info: Info, key: str, valuelen: int = MAX_INFO_VAL) -> Tuple[Optional[str], bool]:
    if key in info.items:
        return info.items[key][:valuelen], True
    return None, False


def Info_get_valuelen(info: Info, key: str) -> Tuple[int, bool]:
    return (len(info.items[key]), True) if key in info.items else (0, False)


def Info_get_nkeys(info: Info) -> int:
    return len(info.items)


def Info_get_nthkey(info: Info, n: int) -> str:
    keys = list(info.items)
    if not 0 <= n < len(keys):
        raise _err(ERR_ARG, "Info_get_nthkey")
    return keys[n]


def Info_delete(info: Info, key: str) -> None:
    if key not in info.items:
        raise _err(ERR_INFO_NOKEY, "Info_delete")
    del info.items[key]


def Info_dup(info: Info) -> Info:
    return Info(info.items)


def Info_free(info: Info) -> None:
    info.items = {}
    return INFO_NULL


# ---- generalized requests, status setters --------------------------------------------------------------------------
class GeneralizedRequest:
    """``Grequest_start``: a request whose completion the application signals with ``Grequest_complete``.
    ``query_fn(extra_state, status)`` fills the status when the request is tested / waited on,
    ``free_fn(extra_state)`` runs when it is released, ``cancel_fn(extra_state, complete)`` on ``Cancel``."""

    def __init__(self, query_fn, free_fn, cancel_fn, extra_state):
        import threading
        self.query_fn, self.free_fn, self.cancel_fn, self.extra_state = query_fn, free_fn, cancel_fn, extra_state
        self._done = threading.Event()
        self._released = False

    # the slice of the Request interface the completion calls use
    def __len__(self):
        return 1

    def is_null(self, i=0):
        return self._released

    def _complete_into(self, status) -> None:
        if self.query_fn:
            self.query_fn(self.extra_state, status)
        if self.free_fn:
            self.free_fn(self.extra_state)
        self._released = True

    def test(self, status) -> bool:
        if self._released:
            return True
        if not self._done.is_set():
            return False
        self._complete_into(status)
        return True

    def wait(self, status) -> None:
        if not self._released:
            self._done.wait()
            self._complete_into(status)

    def cancel(self) -> None:
        if self.cancel_fn:
            self.cancel_fn(self.extra_state, self._done.is_set())


def Grequest_start(query_fn=None, free_fn=None, cancel_fn=None, extra_state=None) -> GeneralizedRequest:
    return GeneralizedRequest(query_fn, free_fn, cancel_fn, extra_state)


def Grequest_complete(request: GeneralizedRequest) -> None:
    request._done.set()


def Status_set_cancelled(status, flag: bool) -> None:
    status[0].cancelled = 1 if flag else 0


def Status_set_elements(status, dtype: int, count: int) -> None:
    from . import datatypes
    t = datatypes.get(dtype)
    sizes = t.elem_sizes()
    full, part = divmod(count, max(1, sizes.size))
    status[0].count_bytes = full * t.size + int(sizes[:part].sum())


def Request_get_status(request, status=None, i: int = 0) -> bool:
    """Completion flag (and status) of a request without releasing it."""
    from .._native import host
    m = _m()
    if isinstance(request, GeneralizedRequest):
        return request._released or request._done.is_set()
    st = status or m.Status()
    f = ctypes.c_int(0)
    m._chk(host().mpit_msg_request_get_status(int(request._arr[i]), ctypes.byref(f), st._ptr()), "Request_get_status")
    return bool(f.value)


# ---- environment / misc --------------------------------------------------------------------------------------------
def Init_thread(required: int = 0) -> int:
    """MPI_Init_thread: initialise and return the provided level (the native layer always provides MULTIPLE)."""
    return _m().Init(thread_level=required)


_PCONTROL = {"level": 1}


def Pcontrol(level: int) -> None:
    """Profiling control: 0 off, 1 default, 2 also flushes.  Drives this framework's timing buckets
    (``mpit_b200.utils``) through ``MPIT_PROFILE_LEVEL``; the message layer's counters are always on."""
    import os
    _PCONTROL["level"] = int(level)
    os.environ["MPIT_PROFILE_LEVEL"] = str(int(level))


_ATTACHED: List[Any] = []


def Buffer_attach(buffer, size: Optional[int] = None) -> None:
    """Provide buffer space for ``Bsend``.  The native Send already copies into the progress thread's queue, so the
    attachment is only recorded (one buffer at a time, as MPI requires)."""
    if _ATTACHED:
        raise _err(ERR_BUFFER, "Buffer_attach: a buffer is already attached")
    _ATTACHED.append(buffer)


def Buffer_detach():
    """-> (buffer, size_in_bytes).  Returns after buffered messages have been handed to the transport."""
    if not _ATTACHED:
        raise _err(ERR_BUFFER, "Buffer_detach: nothing attached")
    b = _ATTACHED.pop()
    n = b.numel() * b.element_size() if isinstance(b, torch.Tensor) else getattr(b, "nbytes", len(b))
    return b, n


def Op_commutative(op: int) -> bool:
    m = _m()
    if m.extras.is_user_op(op) and op >= 100:
        return bool(m.extras._USER_OPS[op][1])
    return True


def Reduce_local(inbuf, inoutbuf, count: int, dtype: int, op: int) -> None:
    """``inoutbuf = inbuf op inoutbuf`` on this process only."""
    m = _m()
    a, b = m.extras._flat(inbuf)[:count], m.extras._flat(inoutbuf)[:count]
    if m.extras.is_user_op(op):
        m.extras._USER_OPS[op][0](a, b)
        return
    if op in (m.SUM, m.PROD, m.MAX, m.MIN):
        m.extras._apply_op(b, a, op)
        return
    is_t = isinstance(b, torch.Tensor)
    if op == m.LAND:
        r = (a != 0) & (b != 0)
    elif op == m.LOR:
        r = (a != 0) | (b != 0)
    elif op == m.BAND:
        r = a & b
    elif op == m.BOR:
        r = a | b
    else:
        raise _err(ERR_OP, "Reduce_local")
    b[:] = r.to(b.dtype) if is_t else r.astype(b.dtype)


def Sendrecv_replace(buf, count: int, dtype: int, dest: int, sendtag: int, source: int, recvtag: int, comm: int = 0,
                     status=None) -> None:
    """Send ``buf`` and receive into the same ``buf``."""
    m = _m()
    flat = m.extras._flat(buf)
    out = flat.clone() if isinstance(flat, torch.Tensor) else flat.copy()
    m.Sendrecv(out, count, dtype, dest, sendtag, buf, count, dtype, source, recvtag, comm, status)


def Alltoallw(sendbuf, sendcounts: Sequence[int], sdispls_bytes: Sequence[int], sendtypes: Sequence[int], recvbuf,
              recvcounts: Sequence[int], rdispls_bytes: Sequence[int], recvtypes: Sequence[int], comm: int = 0) -> None:
    """The most general all-to-all: per-peer counts, BYTE displacements and datatypes (derived types allowed)."""
    from . import datatypes
    m = _m()
    me, n = m.Comm_rank(comm), m.Comm_size(comm)
    sbytes = datatypes._host_bytes(sendbuf, False)
    rbytes = datatypes._host_bytes(recvbuf, True)
    tag = 32100 + 7
    streams = {p: datatypes.pack_bytes(sbytes[sdispls_bytes[p]:], sendcounts[p], sendtypes[p]) for p in range(n)}
    req = m.Request(2 * n)
    incoming = {}
    for p in range(n):
        if p == me:
            continue
        incoming[p] = np.empty(recvcounts[p] * datatypes.get(recvtypes[p]).size, dtype=np.uint8)
        m.Irecv(incoming[p], incoming[p].size, m.BYTE, p, tag, comm, req, p)
        m.Isend(streams[p], streams[p].size, m.BYTE, p, tag, comm, req, n + p)
    incoming[me] = streams[me]
    m.Waitall(req)
    for p in range(n):
        datatypes.unpack_bytes(incoming[p], rbytes[rdispls_bytes[p]:], recvcounts[p], recvtypes[p])


def _expand_ranges(ranges: Sequence[Sequence[int]]) -> List[int]:
    out: List[int] = []
    for first, last, stride in ranges:
        if stride == 0:
            raise _err(ERR_ARG, "Group_range_*: zero stride")
        out.extend(range(first, last + (1 if stride > 0 else -1), stride))
    return out


def Group_range_incl(group, ranges: Sequence[Sequence[int]]):
    """``ranges`` = [(first, last, stride), ...] of ranks in ``group``."""
    return _m().Group_incl(group, _expand_ranges(ranges))


def Group_range_excl(group, ranges: Sequence[Sequence[int]]):
    return _m().Group_excl(group, _expand_ranges(ranges))



# ---- handle arrays (/root/reference/lua-mpi.h:7-21,43-67) ----------------------------------------------------------
class HandleArray(list):
    """The reference exposes ``mpiT.<Type>(n)`` constructors that allocate ``n`` null handles for MPI to write
    through (``mpiT.Request(n)``, ``mpiT.Status(n)``, ``mpiT.Comm(n)`` ...).  Calls here RETURN their results, so only
    ``Request`` and ``Status`` are ever required; the others are plain lists of null handles kept for scripts that
    allocate them up front.  ``Datatype``, ``File``, ``Group``, ``Info`` and ``Win`` are the classes of the objects
    themselves."""

    def __init__(self, kind: str, n: int, null):
        super().__init__([null] * n)
        self.kind = kind


def Aint(n: int = 1) -> HandleArray:
    return HandleArray("Aint", n, 0)


def Offset(n: int = 1) -> HandleArray:
    return HandleArray("Offset", n, 0)


def Comm(n: int = 1) -> HandleArray:
    return HandleArray("Comm", n, -1)                 # COMM_NULL


def Op(n: int = 1) -> HandleArray:
    return HandleArray("Op", n, -1)                   # OP_NULL


def Errhandler(n: int = 1) -> HandleArray:
    return HandleArray("Errhandler", n, -1)           # ERRHANDLER_NULL


__all__ = [_n for _n in dir() if _n[0].isupper() and _n not in ("Any", "Callable", "Dict", "List", "Optional", "Sequence",
                                                                 "Tuple")]
