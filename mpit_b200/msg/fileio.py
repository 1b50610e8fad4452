"""MPI-IO: parallel file access with views, individual / shared file pointers, collective, split-collective and
non-blocking forms.

/root/reference/mpifuncs.c:557-1080 wraps all 52 ``MPI_File_*`` functions ("MPI-IO 52/52"); no
script in the tree calls one -- checkpoints go through ``torch.save``.  They are provided for users who shard large
parameter vectors or datasets across ranks in one file: with a view built from ``Type_create_subarray`` or
``Type_vector`` every rank reads / writes its own slice of a flat fp32 file without a gather on rank 0.

Data movement is native (csrc/msg/hostio.cpp): a view is flattened once into (displacement, length) blocks per tile
and ``mpit_io_view_rw`` walks the tiling with ``pread`` / ``pwrite``; memory-side datatypes go through the pack engine
of ``datatypes.py``.  The shared file pointer lives in a side-car file (``.<name>.shfp``) advanced under an fcntl lock,
so it is shared between processes.  Offsets and pointers are in units of the view's etype, as in MPI.

Conventions: out-parameters are returned; every data-access call accepts an optional ``status`` (``count_bytes`` = bytes
moved) and returns the number of bytes moved; errors raise ``MPIError`` with the MPI-IO error classes.
"""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Callable, Dict, Optional, Tuple

import numpy as np
import torch

from .._native import host
from . import datatypes as dt
from .attrs import GeneralizedRequest, Info

MODE_RDONLY, MODE_RDWR, MODE_WRONLY, MODE_CREATE, MODE_EXCL = 1, 2, 4, 8, 16
MODE_DELETE_ON_CLOSE, MODE_UNIQUE_OPEN, MODE_APPEND, MODE_SEQUENTIAL = 32, 64, 128, 256
SEEK_SET, SEEK_CUR, SEEK_END = 600, 602, 604
DISPLACEMENT_CURRENT = -54278278
FILE_NULL = None
MAX_DATAREP_STRING = 128

(ERR_ACCESS, ERR_AMODE, ERR_BAD_FILE, ERR_CONVERSION, ERR_DUP_DATAREP, ERR_FILE_EXISTS, ERR_FILE_IN_USE, ERR_FILE) = (
    20, 21, 22, 23, 24, 25, 26, 27)
ERR_IO, ERR_NO_SPACE, ERR_NO_SUCH_FILE, ERR_NOT_SAME, ERR_QUOTA, ERR_READ_ONLY = 32, 36, 37, 35, 39, 40
ERR_UNSUPPORTED_DATAREP, ERR_UNSUPPORTED_OPERATION, ERR_ARG, ERR_REQUEST = 43, 44, 13, 7

_DATAREPS: Dict[str, Tuple[Optional[Callable], Optional[Callable], Optional[Callable], object]] = {
    "native": (None, None, None, None), "internal": (None, None, None, None), "external32": (None, None, None, None)}


def _m():
    from .. import msg
    return msg


def _err(code: int, what: str):
    return _m().MPIError(code, what)


def _chk(code: int, what: str):
    if code != 0:
        raise _err(code, what)


def _group(comm: int) -> Tuple[int, int]:
    """(rank, size) of ``comm``; a process that never called Init may still use COMM_SELF files."""
    m = _m()
    if not m.Initialized():
        if comm != m.COMM_SELF:
            raise _err(ERR_ARG, "File_open on a multi-process communicator before Init")
        return 0, 1
    return m.Comm_rank(comm), m.Comm_size(comm)


def _ll(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


class File:
    """An open file (MPI_File)."""

    def __init__(self, path: str, fd: int, amode: int, comm: int, info: Optional[Info], shfd: int):
        self.path, self.fd, self.amode, self.comm, self.shfd = path, fd, amode, comm, shfd
        self.info = Info(info.items) if info is not None else Info()
        self.atomic = False
        self.pos = 0                                   # individual file pointer, in etypes
        self.split: Optional[Tuple[str, int]] = None   # outstanding split collective: (kind, bytes moved)
        self._set_view(0, _m().BYTE, _m().BYTE, "native")

    def _set_view(self, disp: int, etype: int, filetype: int, datarep: str):
        e, f = dt.get(etype), dt.get(filetype)
        if e.size == 0 or f.size % e.size:
            raise _err(ERR_ARG, "File_set_view: the filetype must be built from whole etypes")
        d, n = f.blocks()
        if d.size and (np.any(np.diff(d) < 0) or d.min() < 0):
            raise _err(ERR_ARG, "File_set_view: filetype displacements must be non-negative and non-decreasing")
        self.disp, self.etype, self.filetype, self.datarep = disp, etype, filetype, datarep
        self.esize, self.fsize, self.fextent = e.size, f.size, max(f.extent, 1)
        self.fdisp, self.flen = np.ascontiguousarray(d), np.ascontiguousarray(n)

    # visible bytes between the view start and the end of the file
    def visible_bytes_to(self, file_size: int) -> int:
        span = file_size - self.disp
        if span <= 0:
            return 0
        tiles, rem = divmod(span, self.fextent)
        part = int(np.clip(rem - self.fdisp, 0, self.flen).sum())
        return tiles * self.fsize + part


def _live(fh: File) -> File:
    if fh is None or fh.fd < 0:
        raise _err(ERR_FILE, "invalid file handle")
    return fh


def _sidecar(path: str) -> str:
    d, b = os.path.split(os.path.abspath(path))
    return os.path.join(d, f".{b}.shfp")


# ---- open / close / whole-file operations --------------------------------------------------------------------------
def File_open(comm: int, filename: str, amode: int, info: Optional[Info] = None) -> File:
    """Collective over ``comm``.  With MODE_CREATE (| MODE_EXCL) rank 0 creates the file, everyone then opens it."""
    m = _m()
    access = [bool(amode & b) for b in (MODE_RDONLY, MODE_RDWR, MODE_WRONLY)]
    if sum(access) != 1 or (amode & MODE_RDONLY and amode & (MODE_CREATE | MODE_EXCL)) or \
            (amode & MODE_SEQUENTIAL and amode & MODE_RDWR):
        raise _err(ERR_AMODE, "File_open")
    rank, size = _group(comm)
    h = host()
    fd, shfd = ctypes.c_int(-1), ctypes.c_int(-1)
    rc = 0
    if rank == 0:
        rc = h.mpit_io_open(filename.encode(), amode, ctypes.byref(fd))
        if rc == 0:
            rc = h.mpit_io_open(_sidecar(filename).encode(), MODE_RDWR | MODE_CREATE, ctypes.byref(shfd))
            if rc == 0:
                h.mpit_io_set_size(shfd.value, 0)      # a fresh shared pointer for this open
    if size > 1:
        code = torch.tensor([rc], dtype=torch.int32)
        m.Bcast(code, 1, m.INT, 0, comm)
        rc = int(code[0])
    _chk(rc, f"File_open({filename!r})")
    if rank != 0:
        _chk(h.mpit_io_open(filename.encode(), amode & ~(MODE_EXCL | MODE_CREATE), ctypes.byref(fd)), "File_open")
        _chk(h.mpit_io_open(_sidecar(filename).encode(), MODE_RDWR, ctypes.byref(shfd)), "File_open")
    fh = File(filename, fd.value, amode, comm, info, shfd.value)
    if amode & MODE_APPEND:
        fh.pos = File_get_size(fh) // fh.esize
        if rank == 0:
            _shared(fh, 1, fh.pos)
    if size > 1:
        m.Barrier(comm)
    return fh


def File_close(fh: File) -> None:
    """Collective.  Returns FILE_NULL (None); MODE_DELETE_ON_CLOSE removes the file once everyone has closed it."""
    fh = _live(fh)
    if fh.split is not None:
        raise _err(ERR_REQUEST, "File_close with a split collective outstanding")
    rank, size = _group(fh.comm)
    h = host()
    h.mpit_io_close(fh.fd)
    h.mpit_io_close(fh.shfd)
    fh.fd = fh.shfd = -1
    if size > 1:
        _m().Barrier(fh.comm)
    if rank == 0:
        h.mpit_io_delete(_sidecar(fh.path).encode())
        if fh.amode & MODE_DELETE_ON_CLOSE:
            h.mpit_io_delete(fh.path.encode())
    if size > 1 and fh.amode & MODE_DELETE_ON_CLOSE:
        _m().Barrier(fh.comm)                 # the file is gone on every rank's return
    from . import attrs
    attrs._OBJ_ERRH.pop(attrs._obj_key("file", fh), None)
    return FILE_NULL


def File_delete(filename: str, info: Optional[Info] = None) -> None:
    _chk(host().mpit_io_delete(filename.encode()), f"File_delete({filename!r})")


def File_get_size(fh: File) -> int:
    s = ctypes.c_longlong(0)
    _chk(host().mpit_io_get_size(_live(fh).fd, ctypes.byref(s)), "File_get_size")
    return s.value


def _collective_by_root(fh: File, fn: Callable[[], int], what: str) -> None:
    """Resize-style collectives: rank 0 acts, everyone learns the outcome."""
    m = _m()
    rank, size = _group(fh.comm)
    rc = fn() if rank == 0 else 0
    if size > 1:
        code = torch.tensor([rc], dtype=torch.int32)
        m.Bcast(code, 1, m.INT, 0, fh.comm)
        rc = int(code[0])
    _chk(rc, what)


def File_set_size(fh: File, size: int) -> None:
    fh = _live(fh)
    _writable(fh, "File_set_size")
    _collective_by_root(fh, lambda: host().mpit_io_set_size(fh.fd, size), "File_set_size")


def File_preallocate(fh: File, size: int) -> None:
    fh = _live(fh)
    _writable(fh, "File_preallocate")
    _collective_by_root(fh, lambda: host().mpit_io_preallocate(fh.fd, size), "File_preallocate")


def File_get_group(fh: File):
    m = _m()
    return m.Comm_group(_live(fh).comm) if m.Initialized() else m.Group([0])


def File_get_amode(fh: File) -> int:
    return _live(fh).amode


def File_set_info(fh: File, info: Info) -> None:
    _live(fh).info.items.update(info.items)


def File_get_info(fh: File) -> Info:
    return Info(_live(fh).info.items)


def File_set_atomicity(fh: File, flag: bool) -> None:
    _live(fh).atomic = bool(flag)


def File_get_atomicity(fh: File) -> bool:
    return _live(fh).atomic


def File_sync(fh: File) -> None:
    fh = _live(fh)
    _chk(host().mpit_io_sync(fh.fd), "File_sync")
    if _group(fh.comm)[1] > 1:
        _m().Barrier(fh.comm)


# ---- views ---------------------------------------------------------------------------------------------------------
def File_set_view(fh: File, disp: int, etype: int, filetype: int, datarep: str = "native",
                  info: Optional[Info] = None) -> None:
    """Collective.  Resets the individual and the shared file pointer to zero."""
    fh = _live(fh)
    if datarep not in _DATAREPS:
        raise _err(ERR_UNSUPPORTED_DATAREP, f"File_set_view: data representation {datarep!r}")
    if disp == DISPLACEMENT_CURRENT:
        if not fh.amode & MODE_SEQUENTIAL:
            raise _err(ERR_ARG, "DISPLACEMENT_CURRENT needs MODE_SEQUENTIAL")
        disp = File_get_byte_offset(fh, _shared(fh, 2, 0))
    rank, size = _group(fh.comm)
    if size > 1:
        _m().Barrier(fh.comm)             # nobody may still be using the old shared pointer
    fh._set_view(disp, etype, filetype, datarep)
    fh.pos = 0
    if rank == 0:
        _shared(fh, 1, 0)
    if info is not None:
        File_set_info(fh, info)
    if size > 1:
        _m().Barrier(fh.comm)


def File_get_view(fh: File) -> Tuple[int, int, int, str]:
    fh = _live(fh)
    return fh.disp, fh.etype, fh.filetype, fh.datarep


def File_get_type_extent(fh: File, dtype: int) -> int:
    """Extent of ``dtype`` in the file's data representation."""
    _live(fh)
    return dt.get(dtype).extent


def File_get_byte_offset(fh: File, offset: int) -> int:
    """Absolute byte position of view offset ``offset`` (in etypes)."""
    fh = _live(fh)
    out = ctypes.c_longlong(0)
    _chk(host().mpit_io_view_byte_offset(fh.disp, fh.fextent, fh.fdisp.size, _ll(fh.fdisp), _ll(fh.flen), fh.fsize,
                                         offset * fh.esize, ctypes.byref(out)), "File_get_byte_offset")
    return out.value


def Register_datarep(datarep: str, read_conversion_fn: Optional[Callable], write_conversion_fn: Optional[Callable],
                     dtype_file_extent_fn: Optional[Callable] = None, extra_state=None) -> None:
    """A user data representation.  ``read_conversion_fn(stream: np.ndarray[uint8], dtype, count, extra_state)`` maps
    file bytes to native bytes after a read; ``write_conversion_fn`` the other way before a write.  Both must keep the
    length (representations that change element sizes are not supported)."""
    if datarep in _DATAREPS:
        raise _err(ERR_DUP_DATAREP, f"Register_datarep({datarep!r})")
    if len(datarep) > MAX_DATAREP_STRING:
        raise _err(ERR_ARG, "Register_datarep: name too long")
    _DATAREPS[datarep] = (read_conversion_fn, write_conversion_fn, dtype_file_extent_fn, extra_state)


def _convert(fh: File, stream: np.ndarray, dtype: int, count: int, reading: bool) -> np.ndarray:
    if fh.datarep in ("native", "internal"):
        return stream
    if fh.datarep == "external32":
        t = dt.get(dtype)
        if np.any(t.code == 4):
            raise _err(ERR_CONVERSION, "external32 stores MPI_LONG in four bytes; use MPI_LONG_LONG or MPI_INT in files")
        sizes = np.tile(t.elem_sizes(), count)
        n = min(int(sizes.sum()), stream.size)
        starts = np.cumsum(sizes) - sizes
        out = stream.copy()
        for s in np.unique(sizes):
            if s == 1:
                continue
            st = starts[(sizes == s) & (starts + s <= n)]
            idx = st[:, None] + np.arange(s)[None, :]
            out[idx] = stream[idx[:, ::-1]]
        return out
    rd, wr, _, state = _DATAREPS[fh.datarep]
    fn = rd if reading else wr
    return stream if fn is None else np.ascontiguousarray(fn(stream, dtype, count, state), dtype=np.uint8)


# ---- the one data-access routine everything funnels into -----------------------------------------------------------
def _writable(fh: File, what: str):
    if fh.amode & MODE_RDONLY:
        raise _err(ERR_READ_ONLY, what)


def _access(fh: File, write: bool, pos_etypes: int, buf, count: int, dtype: int, status, what: str) -> int:
    fh = _live(fh)
    if write:
        _writable(fh, what)
    elif fh.amode & MODE_WRONLY:
        raise _err(ERR_ACCESS, f"{what}: the file was opened write-only")
    if pos_etypes < 0:
        raise _err(ERR_ARG, f"{what}: negative offset")
    t = dt.get(dtype)
    nbytes = count * t.size
    if nbytes % fh.esize:
        raise _err(ERR_ARG, f"{what}: the transfer is not a whole number of etypes")
    direct = (not dt.is_derived(dtype) and fh.datarep in ("native", "internal") and
              not (isinstance(buf, torch.Tensor) and buf.is_cuda))
    if write:
        stream = dt._host_bytes(buf, False)[:nbytes] if direct else \
            _convert(fh, dt.pack_bytes(buf, count, dtype), dtype, count, False)
        if stream.size < nbytes:
            raise _err(ERR_ARG, f"{what}: the buffer holds fewer than count elements")
    else:
        stream = dt._host_bytes(buf, True)[:nbytes] if direct else np.empty(nbytes, dtype=np.uint8)
        if stream.size < nbytes:
            raise _err(ERR_ARG, f"{what}: the buffer holds fewer than count elements")
    stream = stream if stream.flags["C_CONTIGUOUS"] else np.ascontiguousarray(stream)
    h = host()
    moved = ctypes.c_longlong(0)
    if fh.atomic:
        h.mpit_io_lock(fh.fd, 2 if write else 1)
    try:
        rc = h.mpit_io_view_rw(fh.fd, int(write), _ll(stream), nbytes, fh.disp, fh.fextent, fh.fdisp.size,
                               _ll(fh.fdisp), _ll(fh.flen), fh.fsize, pos_etypes * fh.esize, ctypes.byref(moved))
    finally:
        if fh.atomic:
            h.mpit_io_lock(fh.fd, 0)
    _chk(rc, what)
    if not write and not direct and moved.value:
        got = _convert(fh, stream[:moved.value], dtype, count, True)
        dt.unpack_bytes(got, buf, count, dtype, limit=moved.value)
    if status is not None:
        status[0].count_bytes = moved.value
        status[0].error = 0
        status[0].cancelled = 0
    return moved.value


def _no_sequential(fh: File, what: str):
    if _live(fh).amode & MODE_SEQUENTIAL:
        raise _err(ERR_UNSUPPORTED_OPERATION, f"{what} on a MODE_SEQUENTIAL file (use the shared file pointer)")


def _etypes(fh: File, nbytes: int) -> int:
    return nbytes // fh.esize


# explicit offsets
def File_read_at(fh: File, offset: int, buf, count: int, dtype: int, status=None) -> int:
    _no_sequential(fh, "File_read_at")
    return _access(fh, False, offset, buf, count, dtype, status, "File_read_at")


def File_write_at(fh: File, offset: int, buf, count: int, dtype: int, status=None) -> int:
    _no_sequential(fh, "File_write_at")
    return _access(fh, True, offset, buf, count, dtype, status, "File_write_at")


# Collective forms: every process of the file's group must call them; no process depends on another's data, so each
# simply performs its own access.
File_read_at_all = File_read_at
File_write_at_all = File_write_at


# individual file pointer
def File_read(fh: File, buf, count: int, dtype: int, status=None) -> int:
    _no_sequential(fh, "File_read")
    n = _access(fh, False, fh.pos, buf, count, dtype, status, "File_read")
    fh.pos += _etypes(fh, n)
    return n


def File_write(fh: File, buf, count: int, dtype: int, status=None) -> int:
    _no_sequential(fh, "File_write")
    n = _access(fh, True, fh.pos, buf, count, dtype, status, "File_write")
    fh.pos += _etypes(fh, n)
    return n


File_read_all = File_read
File_write_all = File_write


def File_seek(fh: File, offset: int, whence: int = SEEK_SET) -> None:
    _no_sequential(fh, "File_seek")
    fh.pos = _seek_target(fh, fh.pos, offset, whence)


def _seek_target(fh: File, cur: int, offset: int, whence: int) -> int:
    if whence == SEEK_SET:
        new = offset
    elif whence == SEEK_CUR:
        new = cur + offset
    elif whence == SEEK_END:
        new = fh.visible_bytes_to(File_get_size(fh)) // fh.esize + offset
    else:
        raise _err(ERR_ARG, "File_seek: whence")
    if new < 0:
        raise _err(ERR_ARG, "File_seek: the pointer would move before the start of the view")
    return new


def File_get_position(fh: File) -> int:
    return _live(fh).pos


# shared file pointer
def _shared(fh: File, op: int, value: int) -> int:
    old = ctypes.c_longlong(0)
    _chk(host().mpit_io_shared_pointer(fh.shfd, op, value, ctypes.byref(old)), "shared file pointer")
    return old.value


def File_read_shared(fh: File, buf, count: int, dtype: int, status=None) -> int:
    fh = _live(fh)
    start = _shared(fh, 0, count * dt.get(dtype).size // fh.esize)
    return _access(fh, False, start, buf, count, dtype, status, "File_read_shared")


def File_write_shared(fh: File, buf, count: int, dtype: int, status=None) -> int:
    fh = _live(fh)
    _writable(fh, "File_write_shared")           # before the pointer moves
    start = _shared(fh, 0, count * dt.get(dtype).size // fh.esize)
    return _access(fh, True, start, buf, count, dtype, status, "File_write_shared")


def _ordered_start(fh: File, count: int, dtype: int) -> int:
    """Collective: the accesses happen as if executed in rank order starting at the shared pointer."""
    m = _m()
    rank, size = _group(fh.comm)
    mine = count * dt.get(dtype).size // fh.esize
    if size == 1:
        return _shared(fh, 0, mine)
    sizes = torch.zeros(size, dtype=torch.int64)
    m.Allgather(torch.tensor([mine], dtype=torch.int64), 1, m.LONG, sizes, fh.comm)
    base = torch.zeros(1, dtype=torch.int64)
    if rank == 0:
        base[0] = _shared(fh, 0, int(sizes.sum()))
    m.Bcast(base, 1, m.LONG, 0, fh.comm)
    return int(base[0]) + int(sizes[:rank].sum())


def File_read_ordered(fh: File, buf, count: int, dtype: int, status=None) -> int:
    fh = _live(fh)
    return _access(fh, False, _ordered_start(fh, count, dtype), buf, count, dtype, status, "File_read_ordered")


def File_write_ordered(fh: File, buf, count: int, dtype: int, status=None) -> int:
    fh = _live(fh)
    _writable(fh, "File_write_ordered")
    return _access(fh, True, _ordered_start(fh, count, dtype), buf, count, dtype, status, "File_write_ordered")


def File_seek_shared(fh: File, offset: int, whence: int = SEEK_SET) -> None:
    """Collective: everyone passes the same arguments; rank 0 moves the pointer."""
    fh = _live(fh)
    rank, size = _group(fh.comm)
    if size > 1:
        _m().Barrier(fh.comm)
    if rank == 0:
        _shared(fh, 1, _seek_target(fh, _shared(fh, 2, 0), offset, whence))
    if size > 1:
        _m().Barrier(fh.comm)


def File_get_position_shared(fh: File) -> int:
    return _shared(_live(fh), 2, 0)


# ---- non-blocking forms: the access runs on a helper thread and completes a generalized request --------------------
def _nonblocking(fn: Callable, *args) -> GeneralizedRequest:
    box = {}

    def query(_state, status):
        if "error" in box:
            raise box["error"]
        if status is not None:
            status[0].count_bytes = box.get("moved", 0)
            status[0].error = 0
            status[0].cancelled = 0

    req = GeneralizedRequest(query, None, None, None)

    def run():
        try:
            box["moved"] = fn(*args)
        except Exception as e:                      # surfaces from Wait / Test
            box["error"] = e
        req._done.set()

    threading.Thread(target=run, daemon=True).start()
    return req


def File_iread_at(fh: File, offset: int, buf, count: int, dtype: int) -> GeneralizedRequest:
    return _nonblocking(File_read_at, fh, offset, buf, count, dtype)


def File_iwrite_at(fh: File, offset: int, buf, count: int, dtype: int) -> GeneralizedRequest:
    return _nonblocking(File_write_at, fh, offset, buf, count, dtype)


def File_iread(fh: File, buf, count: int, dtype: int) -> GeneralizedRequest:
    """The individual pointer advances at the call, by the amount requested, so later calls do not overlap."""
    _no_sequential(fh, "File_iread")
    start = fh.pos
    fh.pos += count * dt.get(dtype).size // fh.esize
    return _nonblocking(_access, fh, False, start, buf, count, dtype, None, "File_iread")


def File_iwrite(fh: File, buf, count: int, dtype: int) -> GeneralizedRequest:
    _no_sequential(fh, "File_iwrite")
    start = fh.pos
    fh.pos += count * dt.get(dtype).size // fh.esize
    return _nonblocking(_access, fh, True, start, buf, count, dtype, None, "File_iwrite")


def File_iread_shared(fh: File, buf, count: int, dtype: int) -> GeneralizedRequest:
    start = _shared(_live(fh), 0, count * dt.get(dtype).size // fh.esize)
    return _nonblocking(_access, fh, False, start, buf, count, dtype, None, "File_iread_shared")


def File_iwrite_shared(fh: File, buf, count: int, dtype: int) -> GeneralizedRequest:
    _writable(_live(fh), "File_iwrite_shared")
    start = _shared(_live(fh), 0, count * dt.get(dtype).size // fh.esize)
    return _nonblocking(_access, fh, True, start, buf, count, dtype, None, "File_iwrite_shared")


# ---- split collectives: begin performs the access, end hands back its status ---------------------------------------
def _begin(fh: File, kind: str, moved: int) -> None:
    fh.split = (kind, moved)


def _split_guard(fh: File, what: str):
    if _live(fh).split is not None:
        raise _err(ERR_REQUEST, f"{what}: one split collective per file handle may be outstanding")


def _end(fh: File, kind: str, status) -> int:
    fh = _live(fh)
    if fh.split is None or fh.split[0] != kind:
        raise _err(ERR_REQUEST, f"File_{kind}_end without the matching begin")
    moved = fh.split[1]
    fh.split = None
    if status is not None:
        status[0].count_bytes = moved
        status[0].error = 0
        status[0].cancelled = 0
    return moved


def File_read_at_all_begin(fh: File, offset: int, buf, count: int, dtype: int) -> None:
    _split_guard(fh, "File_read_at_all_begin")
    _begin(fh, "read_at_all", File_read_at(fh, offset, buf, count, dtype))


def File_read_at_all_end(fh: File, buf=None, status=None) -> int:
    return _end(fh, "read_at_all", status)


def File_write_at_all_begin(fh: File, offset: int, buf, count: int, dtype: int) -> None:
    _split_guard(fh, "File_write_at_all_begin")
    _begin(fh, "write_at_all", File_write_at(fh, offset, buf, count, dtype))


def File_write_at_all_end(fh: File, buf=None, status=None) -> int:
    return _end(fh, "write_at_all", status)


def File_read_all_begin(fh: File, buf, count: int, dtype: int) -> None:
    _split_guard(fh, "File_read_all_begin")
    _begin(fh, "read_all", File_read(fh, buf, count, dtype))


def File_read_all_end(fh: File, buf=None, status=None) -> int:
    return _end(fh, "read_all", status)


def File_write_all_begin(fh: File, buf, count: int, dtype: int) -> None:
    _split_guard(fh, "File_write_all_begin")
    _begin(fh, "write_all", File_write(fh, buf, count, dtype))


def File_write_all_end(fh: File, buf=None, status=None) -> int:
    return _end(fh, "write_all", status)


def File_read_ordered_begin(fh: File, buf, count: int, dtype: int) -> None:
    _split_guard(fh, "File_read_ordered_begin")
    _begin(fh, "read_ordered", File_read_ordered(fh, buf, count, dtype))


def File_read_ordered_end(fh: File, buf=None, status=None) -> int:
    return _end(fh, "read_ordered", status)


def File_write_ordered_begin(fh: File, buf, count: int, dtype: int) -> None:
    _split_guard(fh, "File_write_ordered_begin")
    _begin(fh, "write_ordered", File_write_ordered(fh, buf, count, dtype))


def File_write_ordered_end(fh: File, buf=None, status=None) -> int:
    return _end(fh, "write_ordered", status)


__all__ = [_n for _n in dir() if _n[0].isupper() and _n not in ("Any", "Callable", "Dict", "List", "Optional", "Sequence",
                                                                 "Tuple")]
