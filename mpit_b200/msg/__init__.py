"""Tensor message layer: the ``mpiT`` table of the reference, backed by ``libmpit_host.so``.

Reference surface: /root/reference/lua-mpi.h:235-297 registers 262 MPI wrappers plus handle
constructors and constants as the Lua table ``mpiT``; the tree only ever uses ``Init, Finalize,
Comm_rank, Comm_size, Send, Recv, Isend, Irecv, Iprobe, Test, Wait, Cancel, Test_cancelled,
Allreduce, Iallreduce, Request, Status`` and ``FLOAT, LONG, BYTE, SUM, COMM_WORLD``.
This module provides those, plus ``Bcast, Barrier, Probe, Sendrecv, Reduce, Gather, Allgather,
Scatter, Alltoall, Scan, Reduce_scatter_block, Waitall, Testall, Waitany, Wtime, Wtick, Comm_dup,
Comm_split, Comm_free, Get_count, Get_processor_name, Abort, Init_MTF/MTS/MTM`` natively, and re-exports
the remaining chapters of the binding from its sibling modules: ``extras`` (send modes, persistent requests,
vector collectives, groups, user ops, fence windows), ``datatypes`` (derived datatypes, Pack/Unpack),
``fileio`` (MPI-IO), ``topo`` (cartesian / graph topologies), ``attrs`` (keyvals, error handlers, Info,
generalized requests), ``rma`` (post/start/complete/wait) and ``dynproc`` (inter-communicators, ports,
name service).  ``tools/mpi_coverage.py`` checks the result against the reference's 270 spec files.

Buffers are torch tensors or numpy arrays.  Where the reference passes a Torch *Storage* and reads
``storage->data`` with no type or bounds check (/root/reference/lua-mpi.h:70-76), this layer checks
dtype size and element count.  CUDA tensors are staged through pinned host memory (the reference
hands device pointers to a CUDA-aware MPI, lua-mpi.h:78); the device-to-device fast path of this
framework is the symmetric-heap transport in ``mpit_b200.parallel``, not this layer.

Functions return what MPI writes through out-parameters (``Comm_rank -> int``, ``Test -> bool``...)
and raise :class:`MPIError` on a non-zero error code, instead of returning the code as a number
(/root/reference/mpifuncs.c: every wrapper ends in ``lua_pushnumber(L, res)``).
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import numpy as np
import torch

from .._native import host

# ---- constants (/root/reference/lua-mpi.h:125-231) -------------------------------------------------
BYTE, CHAR, SHORT, INT, LONG, FLOAT, DOUBLE, UNSIGNED_CHAR, UNSIGNED, LONG_LONG = range(10)
SUM, PROD, MAX, MIN, LAND, LOR, BAND, BOR = range(8)
COMM_WORLD, COMM_SELF, COMM_NULL = 0, 1, -1
ANY_SOURCE, ANY_TAG, REQUEST_NULL, UNDEFINED = -1, -1, -1, -32766
SUCCESS = 0
ERR_TRUNCATE = 15
THREAD_SINGLE, THREAD_FUNNELED, THREAD_SERIALIZED, THREAD_MULTIPLE = 0, 1, 2, 3

_DTYPE_SIZE = {BYTE: 1, CHAR: 1, SHORT: 2, INT: 4, LONG: 8, FLOAT: 4, DOUBLE: 8, UNSIGNED_CHAR: 1, UNSIGNED: 4,
               LONG_LONG: 8}
_TORCH_TO_MPI = {torch.float32: FLOAT, torch.float64: DOUBLE, torch.int64: LONG, torch.int32: INT,
                 torch.int16: SHORT, torch.uint8: BYTE, torch.int8: CHAR}

# signal constants live with the coroutine layer but the reference hangs them on the same table
# (/root/reference/init.lua:21-25)
signal_INIT, signal_EXEC, signal_OK, signal_ERR, signal_DONE = 1, 2, 3, 4, 5


class MPIError(RuntimeError):
    def __init__(self, code: int, what: str):
        super().__init__(f"{what} failed with error class {code}")
        self.code = code


class _CStatus(ctypes.Structure):
    _fields_ = [("source", ctypes.c_int), ("tag", ctypes.c_int), ("error", ctypes.c_int),
                ("count_bytes", ctypes.c_longlong), ("cancelled", ctypes.c_int)]


class Status:
    """``mpiT.Status(n)`` (/root/reference/lua-mpi.h:7-21): an array of n status records."""

    def __init__(self, n: int = 1):
        self._arr = (_CStatus * n)()

    def __len__(self):
        return len(self._arr)

    def __getitem__(self, i):
        return self._arr[i]

    source = property(lambda s: s._arr[0].source)
    tag = property(lambda s: s._arr[0].tag)
    error = property(lambda s: s._arr[0].error)
    count_bytes = property(lambda s: s._arr[0].count_bytes)
    cancelled = property(lambda s: bool(s._arr[0].cancelled))

    def _ptr(self, i=0):
        return ctypes.byref(self._arr[i])


class Request:
    """``mpiT.Request(n)``: n request handles, initialised to REQUEST_NULL."""

    def __init__(self, n: int = 1):
        self._arr = (ctypes.c_int * n)(*([REQUEST_NULL] * n))
        self._after: List[Optional[callable]] = [None] * n   # completion hooks (device staging)
        self._keep: List[object] = [None] * n                # buffers that must outlive the transfer
        self._unroute: List[int] = [0] * n                   # inter-communicator source offset of a receive
        self._dev: List[object] = [None] * n                 # live request of the device transport (msg/device.py)

    def __len__(self):
        return len(self._arr)

    def is_null(self, i=0):
        return self._arr[i] == REQUEST_NULL

    def _ptr(self, i=0):
        return ctypes.byref(self._arr, i * ctypes.sizeof(ctypes.c_int))

    def _finish(self, i=0):
        fn, self._after[i], self._keep[i] = self._after[i], None, None
        if fn is not None:
            fn()


def _chk(code: int, what: str):
    if code != SUCCESS:
        raise MPIError(code, what)


def _intra(comm, what: str):
    """Collectives take intra-communicators; an inter-communicator has to be merged first (Intercomm_merge)."""
    if not isinstance(comm, int):
        raise MPIError(5, f"{what} on an inter-communicator (use Intercomm_merge and call it on the result)")
    return comm


class _Buf:
    """Resolve a user buffer to (host pointer, keep-alive, completion hook)."""

    def __init__(self, obj, count: int, dtype: int, writable: bool):
        self.after = None
        self.count, self.dtype = count, dtype          # what the native layer is told (bytes for a derived type)
        if dtype >= datatypes.DERIVED_BASE:
            stream, self.after = datatypes.stage_for_message(obj, count, dtype, writable)
            self.ptr, self.keep = ctypes.c_void_p(stream.ctypes.data), (obj, stream)
            self.count, self.dtype = stream.size, BYTE
            return
        es = _DTYPE_SIZE[dtype]
        nbytes = count * es
        if obj is None:
            if count != 0:
                raise ValueError("a null buffer can only carry zero elements")
            self.ptr, self.keep = ctypes.c_void_p(0), None
            return
        if isinstance(obj, torch.Tensor):
            if not obj.is_contiguous():
                raise ValueError("message buffers must be contiguous")
            if dtype not in (BYTE, CHAR, UNSIGNED_CHAR) and obj.element_size() != es:
                raise TypeError(f"tensor element size {obj.element_size()} does not match the MPI datatype ({es} bytes)")
            if obj.numel() * obj.element_size() < nbytes:
                raise ValueError(f"buffer holds {obj.numel()} elements, {count} requested")
            if obj.is_cuda:
                flat = obj.view(-1)
                n_el = (nbytes + obj.element_size() - 1) // obj.element_size()
                stage = torch.empty(n_el, dtype=obj.dtype, pin_memory=True)
                if writable:
                    self.after = lambda: flat[:n_el].copy_(stage)
                else:
                    stage.copy_(flat[:n_el])
                    torch.cuda.current_stream(obj.device).synchronize()
                self.ptr, self.keep = ctypes.c_void_p(stage.data_ptr()), (obj, stage)
            else:
                self.ptr, self.keep = ctypes.c_void_p(obj.data_ptr()), obj
            return
        if isinstance(obj, np.ndarray):
            if not obj.flags["C_CONTIGUOUS"]:
                raise ValueError("message buffers must be contiguous")
            if obj.nbytes < nbytes:
                raise ValueError("buffer too small")
            self.ptr, self.keep = ctypes.c_void_p(obj.ctypes.data), obj
            return
        if isinstance(obj, (bytearray, memoryview)):
            arr = (ctypes.c_char * len(obj)).from_buffer(obj)
            if len(obj) < nbytes:
                raise ValueError("buffer too small")
            self.ptr, self.keep = ctypes.cast(arr, ctypes.c_void_p), (obj, arr)
            return
        if isinstance(obj, bytes) and not writable:
            self.ptr, self.keep = ctypes.cast(ctypes.c_char_p(obj), ctypes.c_void_p), obj
            return
        raise TypeError(f"unsupported buffer type {type(obj)}")


def mpi_dtype_of(t: torch.Tensor) -> int:
    """The reference's stubbed ``mpiT.type(s)`` (/root/reference/init.lua:121-123), implemented."""
    return _TORCH_TO_MPI[t.dtype]


# ---- environment ---------------------------------------------------------------------------------------
_state = {"rank": 0, "size": 1}


def Init(rank: Optional[int] = None, size: Optional[int] = None, master_addr: Optional[str] = None,
         master_port: Optional[int] = None, thread_level: int = THREAD_SINGLE) -> int:
    """MPI_Init (/root/reference/lua-mpi.h:80-85).  Rank/size/rendezvous come from the environment
    (``MPIT_RANK``/``RANK``, ``MPIT_WORLD_SIZE``/``WORLD_SIZE``, ``MASTER_ADDR``, ``MPIT_PORT`` or
    ``MASTER_PORT``+17) as set by ``python -m mpit_b200.launch.run`` or torchrun."""
    env = os.environ
    rank = int(env.get("MPIT_RANK", env.get("RANK", "0"))) if rank is None else rank
    size = int(env.get("MPIT_WORLD_SIZE", env.get("WORLD_SIZE", "1"))) if size is None else size
    master_addr = master_addr or env.get("MPIT_ADDR", env.get("MASTER_ADDR", "127.0.0.1"))
    if master_port is None:
        master_port = int(env["MPIT_PORT"]) if "MPIT_PORT" in env else int(env.get("MASTER_PORT", "29400")) + 17
    provided = ctypes.c_int(0)
    _chk(host().mpit_msg_init(rank, size, master_addr.encode(), master_port, thread_level, ctypes.byref(provided)),
         "Init")
    if provided.value < thread_level:
        Abort(COMM_WORLD, 1)
    _state["rank"], _state["size"] = rank, size
    return provided.value


def Init_MTF():
    """MPI_Init_thread(FUNNELED) (/root/reference/lua-mpi.h:87-97)."""
    return Init(thread_level=THREAD_FUNNELED)


def Init_MTS():
    """MPI_Init_thread(SERIALIZED) (/root/reference/lua-mpi.h:99-109)."""
    return Init(thread_level=THREAD_SERIALIZED)


def Init_MTM():
    """MPI_Init_thread(MULTIPLE) (/root/reference/lua-mpi.h:111-121)."""
    return Init(thread_level=THREAD_MULTIPLE)


def Initialized() -> bool:
    f = ctypes.c_int(0)
    host().mpit_msg_initialized(ctypes.byref(f))
    return bool(f.value)


def Finalize():
    _chk(host().mpit_msg_finalize(), "Finalize")


def Abort(comm: int = COMM_WORLD, errorcode: int = 1):
    host().mpit_msg_abort(comm, errorcode)


def Wtime() -> float:
    return host().mpit_msg_wtime()


def Wtick() -> float:
    return host().mpit_msg_wtick()


def Get_processor_name() -> str:
    buf = ctypes.create_string_buffer(64)
    n = ctypes.c_int(0)
    _chk(host().mpit_msg_get_processor_name(buf, ctypes.byref(n)), "Get_processor_name")
    return buf.value.decode()


def Comm_rank(comm: int = COMM_WORLD) -> int:
    """Rank in ``comm`` (in the LOCAL group of an inter-communicator)."""
    comm = getattr(comm, "local_comm", comm)
    r = ctypes.c_int(0)
    _chk(host().mpit_msg_comm_rank(comm, ctypes.byref(r)), "Comm_rank")
    return r.value


def Comm_size(comm: int = COMM_WORLD) -> int:
    comm = getattr(comm, "local_comm", comm)
    r = ctypes.c_int(0)
    _chk(host().mpit_msg_comm_size(comm, ctypes.byref(r)), "Comm_size")
    return r.value


# helpers of /root/reference/init.lua:28-38
get_rank = Comm_rank
get_size = Comm_size


def Comm_dup(comm: int) -> int:
    """Duplicates the communicator together with its cached attributes (through their copy callbacks) and its
    topology."""
    comm = _intra(comm, "Comm_dup")
    r = ctypes.c_int(0)
    _chk(host().mpit_msg_comm_dup(comm, ctypes.byref(r)), "Comm_dup")
    attrs.copy_attrs("comm", comm, r.value)
    if comm in topo._TOPO:
        topo._TOPO[r.value] = dict(topo._TOPO[comm])
    return r.value


def Comm_split(comm: int, color: int, key: int) -> int:
    comm = _intra(comm, "Comm_split")
    r = ctypes.c_int(0)
    _chk(host().mpit_msg_comm_split(comm, color, key, ctypes.byref(r)), "Comm_split")
    return r.value


def Comm_free(comm) -> None:
    if isinstance(comm, dynproc.Intercomm):
        dynproc.Comm_disconnect(comm)
        return
    c = ctypes.c_int(comm)
    _chk(host().mpit_msg_comm_free(ctypes.byref(c)), "Comm_free")
    attrs.drop_object("comm", comm)
    topo.forget(comm)


# ---- point to point ---------------------------------------------------------------------------------------
_DEV_LIVE = -2       # Request._arr marker of a live device-transport request


def _dev_route(buf, count: int, dtype: int, comm):
    """The device transport takes a message when it is enabled, the buffer is a CUDA tensor, the datatype is a
    basic one and the communicator is COMM_WORLD; returns (transport, nbytes) or (None, 0)."""
    from . import device
    tr = device.transport()
    if tr is None or comm != COMM_WORLD or not isinstance(buf, torch.Tensor) or not buf.is_cuda \
            or dtype >= datatypes.DERIVED_BASE:
        return None, 0
    if not buf.is_contiguous():
        raise ValueError("message buffers must be contiguous")
    nbytes = count * _DTYPE_SIZE[dtype]
    if buf.numel() * buf.element_size() < nbytes:
        raise ValueError(f"buffer holds {buf.numel()} elements, {count} requested")
    return tr, nbytes


def _dev_status(status: Optional["Status"], req) -> None:
    if status is not None:
        s = status[0]
        s.source, s.tag, s.error, s.count_bytes, s.cancelled = req.src, req.rtag, req.error, req.count_bytes, int(req.cancelled)


def _dev_finish(request: "Request", i: int, status: Optional["Status"]) -> None:
    req = request._dev[i]
    _dev_status(status, req)
    request._dev[i] = None
    request._arr[i] = REQUEST_NULL
    if req.error and not req.cancelled:
        raise MPIError(req.error, "message truncated: the receive buffer is smaller than the message")


def _unroute(status: Optional["Status"], offset: int) -> None:
    """Sources reported on an inter-communicator are ranks of the remote group."""
    if status is not None and offset and status[0].source >= 0:
        status[0].source -= offset


def Send(buf, count: int, dtype: int, dest: int, tag: int, comm: int = COMM_WORLD) -> None:
    """``dtype`` may be a committed derived datatype (the message travels packed); ``comm`` may be an
    inter-communicator (``dest`` is then a rank of the remote group)."""
    comm, dest, _ = dynproc.route(comm, dest)
    tr, nbytes = _dev_route(buf, count, dtype, comm)
    if tr is not None:
        tr.wait(tr.isend(buf, nbytes, dest, tag))
        return
    b = _Buf(buf, count, dtype, False)
    _chk(host().mpit_msg_send(b.ptr, b.count, b.dtype, dest, tag, comm), "Send")


def Recv(buf, count: int, dtype: int, source: int, tag: int, comm: int = COMM_WORLD,
         status: Optional[Status] = None) -> None:
    comm, source, off = dynproc.route(comm, source)
    tr, nbytes = _dev_route(buf, count, dtype, comm)
    if tr is not None:
        req = tr.irecv(buf, nbytes, source, tag)
        tr.wait(req)
        _dev_status(status, req)
        if req.error:
            raise MPIError(req.error, "message truncated: the receive buffer is smaller than the message")
        return
    b = _Buf(buf, count, dtype, True)
    _chk(host().mpit_msg_recv(b.ptr, b.count, b.dtype, source, tag, comm, status._ptr() if status else None), "Recv")
    _unroute(status, off)
    if b.after:
        b.after()


def Isend(buf, count: int, dtype: int, dest: int, tag: int, comm: int, request: Request, i: int = 0) -> None:
    comm, dest, _ = dynproc.route(comm, dest)
    tr, nbytes = _dev_route(buf, count, dtype, comm)
    if tr is not None:
        request._dev[i], request._keep[i] = tr.isend(buf, nbytes, dest, tag), buf
        request._arr[i] = _DEV_LIVE
        return
    b = _Buf(buf, count, dtype, False)
    request._keep[i] = b.keep
    _chk(host().mpit_msg_isend(b.ptr, b.count, b.dtype, dest, tag, comm, request._ptr(i)), "Isend")


def Irecv(buf, count: int, dtype: int, source: int, tag: int, comm: int, request: Request, i: int = 0) -> None:
    comm, source, off = dynproc.route(comm, source)
    tr, nbytes = _dev_route(buf, count, dtype, comm)
    if tr is not None:
        request._dev[i], request._keep[i] = tr.irecv(buf, nbytes, source, tag), buf
        request._arr[i] = _DEV_LIVE
        return
    b = _Buf(buf, count, dtype, True)
    request._keep[i], request._after[i] = b.keep, b.after
    request._unroute[i] = off
    _chk(host().mpit_msg_irecv(b.ptr, b.count, b.dtype, source, tag, comm, request._ptr(i)), "Irecv")


def Sendrecv(sbuf, scount, sdtype, dest, stag, rbuf, rcount, rdtype, source, rtag, comm=COMM_WORLD,
             status: Optional[Status] = None) -> None:
    native, dest, off = dynproc.route(comm, dest)
    _, source, _ = dynproc.route(comm, source)
    s, r = _Buf(sbuf, scount, sdtype, False), _Buf(rbuf, rcount, rdtype, True)
    _chk(host().mpit_msg_sendrecv(s.ptr, s.count, s.dtype, dest, stag, r.ptr, r.count, r.dtype, source, rtag, native,
                                  status._ptr() if status else None), "Sendrecv")
    _unroute(status, off)
    if r.after:
        r.after()


def Iprobe(source: int, tag: int, comm: int = COMM_WORLD, status: Optional[Status] = None) -> bool:
    comm, source, off = dynproc.route(comm, source)
    f = ctypes.c_int(0)
    _chk(host().mpit_msg_iprobe(source, tag, comm, ctypes.byref(f), status._ptr() if status else None), "Iprobe")
    if f.value:
        _unroute(status, off)
        return True
    from . import device
    tr = device.transport()
    if tr is not None and comm == COMM_WORLD:
        # messages that travel GPU to GPU: the probe is a scan of shared descriptor words, no kernel launch
        flag, src, mtag, nbytes = tr.iprobe(source, tag)
        if flag and status is not None:
            s0 = status[0]
            s0.source, s0.tag, s0.error, s0.count_bytes, s0.cancelled = src, mtag, 0, nbytes, 0
        return flag
    return False


def Probe(source: int, tag: int, comm: int = COMM_WORLD, status: Optional[Status] = None) -> None:
    from . import device
    if device.transport() is not None and comm == COMM_WORLD:
        import time
        while not Iprobe(source, tag, comm, status):     # either transport may deliver it
            time.sleep(0)
        return
    comm, source, off = dynproc.route(comm, source)
    _chk(host().mpit_msg_probe(source, tag, comm, status._ptr() if status else None), "Probe")
    _unroute(status, off)


def Test(request: Request, status: Optional[Status] = None, i: int = 0) -> bool:
    """MPI_Test: True when complete (the handle becomes REQUEST_NULL).  A truncated receive raises."""
    if isinstance(request, attrs.GeneralizedRequest):
        return request.test(status)
    if request._dev[i] is not None:
        from . import device
        if not device.transport().test(request._dev[i]):
            return False
        _dev_finish(request, i, status)
        return True
    st = status or Status()
    f = ctypes.c_int(0)
    code = host().mpit_msg_test(request._ptr(i), ctypes.byref(f), st._ptr())
    if f.value and not st.cancelled:
        _unroute(st, request._unroute[i])
        request._finish(i)
    elif f.value:
        request._after[i] = request._keep[i] = None
    _chk(code, "Test")
    return bool(f.value)


def Wait(request: Request, status: Optional[Status] = None, i: int = 0) -> None:
    if isinstance(request, attrs.GeneralizedRequest):
        return request.wait(status)
    if request._dev[i] is not None:
        from . import device
        device.transport().wait(request._dev[i])
        _dev_finish(request, i, status)
        return
    st = status or Status()
    was_live = not request.is_null(i)
    code = host().mpit_msg_wait(request._ptr(i), st._ptr())
    if not st.cancelled:
        if was_live:
            _unroute(st, request._unroute[i])
        request._finish(i)
    else:
        request._after[i] = request._keep[i] = None
    _chk(code, "Wait")


def Waitall(request: Request, statuses: Optional[Status] = None) -> None:
    for i in range(len(request)):
        if not request.is_null(i):
            st = Status()
            Wait(request, st, i)
            if statuses is not None:
                ctypes.memmove(statuses._ptr(i), st._ptr(), ctypes.sizeof(_CStatus))


def Testall(request: Request) -> bool:
    return all(request.is_null(i) or Test(request, None, i) for i in range(len(request)))


def Waitany(request: Request, status: Optional[Status] = None) -> int:
    import time
    while True:
        live = False
        for i in range(len(request)):
            if request.is_null(i):
                continue
            live = True
            if Test(request, status, i):
                return i
        if not live:
            return UNDEFINED
        time.sleep(0)


def Cancel(request: Request, i: int = 0) -> None:
    if isinstance(request, attrs.GeneralizedRequest):
        return request.cancel()
    if request._dev[i] is not None:
        from . import device
        device.transport().cancel(request._dev[i])
        return
    _chk(host().mpit_msg_cancel(request._ptr(i)), "Cancel")


def Test_cancelled(status: Status) -> bool:
    f = ctypes.c_int(0)
    _chk(host().mpit_msg_test_cancelled(status._ptr(), ctypes.byref(f)), "Test_cancelled")
    return bool(f.value)


def Get_count(status: Status, dtype: int) -> int:
    """Whole elements of ``dtype`` received (UNDEFINED when the byte count is not a multiple of its size)."""
    if dtype >= datatypes.DERIVED_BASE:
        return datatypes.count_from_bytes(int(status[0].count_bytes), dtype)
    c = ctypes.c_longlong(0)
    _chk(host().mpit_msg_get_count(status._ptr(), dtype, ctypes.byref(c)), "Get_count")
    return c.value


# ---- collectives ----------------------------------------------------------------------------------------------
def Barrier(comm: int = COMM_WORLD) -> None:
    comm = _intra(comm, "Barrier")
    _chk(host().mpit_msg_barrier(comm), "Barrier")


def Bcast(buf, count: int, dtype: int, root: int, comm: int = COMM_WORLD) -> None:
    comm = _intra(comm, "Bcast")
    if dtype == FLOAT and isinstance(buf, torch.Tensor) and buf.is_cuda:
        from . import device
        if device.device_bcast(buf, count, root, comm):
            return
    b = _Buf(buf, count, dtype, True)
    if b.after and Comm_rank(comm) == root and dtype < datatypes.DERIVED_BASE:
        b.keep[1].copy_(b.keep[0].view(-1)[:b.keep[1].numel()])   # root's device data must be staged out first
    _chk(host().mpit_msg_bcast(b.ptr, b.count, b.dtype, root, comm), "Bcast")
    if b.after:
        b.after()


def Reduce(sbuf, rbuf, count: int, dtype: int, op: int, root: int, comm: int = COMM_WORLD) -> None:
    comm = _intra(comm, "Reduce")
    if extras.needs_packed_reduce(dtype, op):
        return extras.reduce_packed(sbuf, rbuf, count, dtype, op, root, comm, everyone=False)
    if extras.is_user_op(op):
        return extras.reduce_user_op(sbuf, rbuf, count, dtype, op, root, comm, everyone=False)
    s, r = _Buf(sbuf, count, dtype, False), _Buf(rbuf, count, dtype, True)
    _chk(host().mpit_msg_reduce(s.ptr, r.ptr, count, dtype, op, root, comm), "Reduce")
    if r.after and Comm_rank(comm) == root:
        r.after()


def Allreduce(sbuf, rbuf, count: int, dtype: int, op: int, comm: int = COMM_WORLD) -> None:
    """MPI_Allreduce (/root/reference/test/testreduceall.lua:28).  The ten basic datatypes with SUM..BOR run in the
    native reducer; user operations, LXOR / BXOR, MAXLOC / MINLOC and the packed datatypes are folded on rank 0."""
    comm = _intra(comm, "Allreduce")
    if op == SUM and dtype == FLOAT and isinstance(sbuf, torch.Tensor) and isinstance(rbuf, torch.Tensor) \
            and sbuf.is_cuda and rbuf.is_cuda:
        from . import device
        if device.device_allreduce(sbuf, rbuf, count, comm):      # NVLink / NVLS kernel, no host staging
            return
    if extras.needs_packed_reduce(dtype, op):
        return extras.reduce_packed(sbuf, rbuf, count, dtype, op, 0, comm, everyone=True)
    if extras.is_user_op(op):
        return extras.reduce_user_op(sbuf, rbuf, count, dtype, op, 0, comm, everyone=True)
    s, r = _Buf(sbuf, count, dtype, False), _Buf(rbuf, count, dtype, True)
    _chk(host().mpit_msg_allreduce(s.ptr, r.ptr, count, dtype, op, comm), "Allreduce")
    if r.after:
        r.after()


def Iallreduce(sbuf, rbuf, count: int, dtype: int, op: int, comm: int, request: Request, i: int = 0) -> None:
    """MPI_Iallreduce (/root/reference/test/testireduceall.lua:32); complete it with Test/Wait."""
    comm = _intra(comm, "Iallreduce")
    s, r = _Buf(sbuf, count, dtype, False), _Buf(rbuf, count, dtype, True)
    request._keep[i], request._after[i] = (s.keep, r.keep), r.after
    _chk(host().mpit_msg_iallreduce(s.ptr, r.ptr, count, dtype, op, comm, request._ptr(i)), "Iallreduce")


def Gather(sbuf, scount: int, dtype: int, rbuf, root: int, comm: int = COMM_WORLD) -> None:
    comm = _intra(comm, "Gather")
    n = Comm_size(comm)
    s = _Buf(sbuf, scount, dtype, False)
    r = _Buf(rbuf, scount * n, dtype, True) if Comm_rank(comm) == root else _Buf(None, 0, dtype, True)
    _chk(host().mpit_msg_gather(s.ptr, scount, dtype, r.ptr, root, comm), "Gather")
    if r.after:
        r.after()


def Allgather(sbuf, scount: int, dtype: int, rbuf, comm: int = COMM_WORLD) -> None:
    comm = _intra(comm, "Allgather")
    s, r = _Buf(sbuf, scount, dtype, False), _Buf(rbuf, scount * Comm_size(comm), dtype, True)
    _chk(host().mpit_msg_allgather(s.ptr, scount, dtype, r.ptr, comm), "Allgather")
    if r.after:
        r.after()


def Scatter(sbuf, count: int, dtype: int, rbuf, root: int, comm: int = COMM_WORLD) -> None:
    comm = _intra(comm, "Scatter")
    n = Comm_size(comm)
    s = _Buf(sbuf, count * n, dtype, False) if Comm_rank(comm) == root else _Buf(None, 0, dtype, False)
    r = _Buf(rbuf, count, dtype, True)
    _chk(host().mpit_msg_scatter(s.ptr, count, dtype, r.ptr, root, comm), "Scatter")
    if r.after:
        r.after()


def Alltoall(sbuf, count: int, dtype: int, rbuf, comm: int = COMM_WORLD) -> None:
    comm = _intra(comm, "Alltoall")
    n = Comm_size(comm)
    s, r = _Buf(sbuf, count * n, dtype, False), _Buf(rbuf, count * n, dtype, True)
    _chk(host().mpit_msg_alltoall(s.ptr, count, dtype, r.ptr, comm), "Alltoall")
    if r.after:
        r.after()


def Reduce_scatter_block(sbuf, rbuf, count: int, dtype: int, op: int, comm: int = COMM_WORLD) -> None:
    comm = _intra(comm, "Reduce_scatter_block")
    s, r = _Buf(sbuf, count * Comm_size(comm), dtype, False), _Buf(rbuf, count, dtype, True)
    _chk(host().mpit_msg_reduce_scatter_block(s.ptr, r.ptr, count, dtype, op, comm), "Reduce_scatter_block")
    if r.after:
        r.after()


def Scan(sbuf, rbuf, count: int, dtype: int, op: int, comm: int = COMM_WORLD) -> None:
    comm = _intra(comm, "Scan")
    s, r = _Buf(sbuf, count, dtype, False), _Buf(rbuf, count, dtype, True)
    _chk(host().mpit_msg_scan(s.ptr, r.ptr, count, dtype, op, comm), "Scan")
    if r.after:
        r.after()


def stats() -> dict:
    a = (ctypes.c_longlong * 8)()
    host().mpit_msg_stats(a)
    keys = ("msgs_sent", "bytes_sent", "msgs_recv", "bytes_recv", "unexpected_now", "posted_now", "requests_live",
            "rendezvous_peers")
    return dict(zip(keys, list(a)))


# ---- serialize / deserialize (/root/reference/init.lua:111-132) -----------------------------------------------
def serialize(obj) -> torch.Tensor:
    """Any Torch7-representable object -> uint8 tensor (the reference's CharStorage), using the Torch7
    binary format so that the bytes are interchangeable with ``torch.MemoryFile`` output."""
    from ..t7 import dumps
    return torch.frombuffer(bytearray(dumps(obj)), dtype=torch.uint8)


def deserialize(buf: torch.Tensor):
    from ..t7 import loads
    return loads(bytes(buf.cpu().numpy().tobytes()))


# coroutine IO lives in mpit_b200.aio; re-exported here because the reference hangs everything on mpiT
from ..aio import (aio_send, aio_recv, co_execute, co_ping, co_wait, Queue)  # noqa: E402,F401


# ---- the wider MPI surface (send modes, persistent requests, vector collectives, groups, user ops, fence
# windows, error classes ...), composed from the calls above: see msg/extras.py ----
from . import extras  # noqa: E402
from .extras import *  # noqa: E402,F401,F403
# ---- the remaining chapters: derived datatypes, MPI-IO, topologies, caching / error handlers / Info, PSCW windows,
# inter-communicators and ports ----
from . import attrs, datatypes, dynproc, fileio, rma, topo  # noqa: E402
from .attrs import *  # noqa: E402,F401,F403
from .datatypes import *  # noqa: E402,F401,F403
from .dynproc import *  # noqa: E402,F401,F403
from .fileio import *  # noqa: E402,F401,F403
from .rma import *  # noqa: E402,F401,F403
from .topo import *  # noqa: E402,F401,F403
from .extras import (PROC_NULL, ROOT, IDENT, CONGRUENT, SIMILAR, UNEQUAL, LXOR, BXOR, MAXLOC, MINLOC,  # noqa: E402,F401
                     GROUP_EMPTY)
