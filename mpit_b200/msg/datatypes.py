"""Derived datatypes: the "Datatypes" chapter of the reference's binding.

/root/reference/mpifuncs.c wraps the MPI type constructors (``Type_contiguous``, ``Type_vector``, ``Type_indexed``,
``Type_create_struct``, ``Type_create_subarray``, ``Type_create_darray`` ...), ``Pack`` / ``Unpack`` and their
``external32`` forms, and the type queries ("Datatypes 26/26"); no script in the tree calls
them.  Here a datatype is a *type map* -- for every basic element its byte displacement and basic type -- plus lower
bound and extent.  Committing a type flattens the map into maximal contiguous byte blocks; the data movement between
a strided user buffer and a contiguous stream is then one call into the native engine
(``mpit_dt_pack`` / ``mpit_dt_unpack``, csrc/msg/hostio.cpp).  CUDA tensors are gathered / scattered on the device
by ``mpit_dt_move_device`` (csrc/kernels/dtpack.cu), so one contiguous copy crosses PCIe.

A committed handle can be given to ``Send / Recv / Isend / Irecv / Sendrecv / Bcast`` and to the ``File_*``
routines wherever a basic datatype is accepted: the message travels packed, as bytes.

Conventions follow the rest of ``mpit_b200.msg``: values MPI returns through out-parameters are returned, handles are
small integers (basic types 0..9, derived types from 1000 up), errors raise ``MPIError``.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .._native import host

DERIVED_BASE = 1000

# combiners (Type_get_envelope)
(COMBINER_NAMED, COMBINER_DUP, COMBINER_CONTIGUOUS, COMBINER_VECTOR, COMBINER_HVECTOR, COMBINER_INDEXED,
 COMBINER_HINDEXED, COMBINER_INDEXED_BLOCK, COMBINER_STRUCT, COMBINER_SUBARRAY, COMBINER_DARRAY,
 COMBINER_RESIZED) = range(1, 13)
ORDER_C, ORDER_FORTRAN = 56, 57
DISTRIBUTE_BLOCK, DISTRIBUTE_CYCLIC, DISTRIBUTE_NONE, DISTRIBUTE_DFLT_DARG = 121, 122, 123, -49767
TYPECLASS_REAL, TYPECLASS_INTEGER, TYPECLASS_COMPLEX = 1, 2, 3
BOTTOM = 0

ERR_TYPE, ERR_ARG, ERR_TRUNCATE, ERR_UNSUPPORTED_DATAREP = 3, 13, 15, 43

# basic types: numpy dtype in memory, and size / dtype in the "external32" representation (big endian,
# MPI_LONG is four bytes there)
# (codes 10..12 are the element kinds of the predefined types the native layer has no wire id for:
# UNSIGNED_SHORT, UNSIGNED_LONG, LONG_DOUBLE -- they travel packed, like derived types)
_NP = {0: np.uint8, 1: np.int8, 2: np.int16, 3: np.int32, 4: np.int64, 5: np.float32, 6: np.float64, 7: np.uint8,
       8: np.uint32, 9: np.int64, 10: np.uint16, 11: np.uint64, 12: np.longdouble}
_EXT32 = {0: ">u1", 1: ">i1", 2: ">i2", 3: ">i4", 4: ">i4", 5: ">f4", 6: ">f8", 7: ">u1", 8: ">u4", 9: ">i8",
          10: ">u2", 11: ">u8", 12: ">f16"}
_NKINDS = len(_NP)
_BASIC_NAMES = {0: "MPI_BYTE", 1: "MPI_CHAR", 2: "MPI_SHORT", 3: "MPI_INT", 4: "MPI_LONG", 5: "MPI_FLOAT",
                6: "MPI_DOUBLE", 7: "MPI_UNSIGNED_CHAR", 8: "MPI_UNSIGNED", 9: "MPI_LONG_LONG"}


def _err(code: int, what: str):
    from . import MPIError
    return MPIError(code, what)


class Datatype:
    """Type map + bounds.  ``disp``/``code`` list the basic elements in type-map order."""

    def __init__(self, disp: np.ndarray, code: np.ndarray, lb: Optional[int] = None, extent: Optional[int] = None,
                 combiner: int = COMBINER_NAMED, contents: Tuple[list, list, list] = ((), (), ())):
        self.disp = np.asarray(disp, dtype=np.int64).reshape(-1)
        self.code = np.asarray(code, dtype=np.int16).reshape(-1)
        sizes = self.elem_sizes()
        self.size = int(sizes.sum())
        if self.disp.size:
            self.true_lb = int(self.disp.min())
            self.true_ub = int((self.disp + sizes).max())
            align = int(sizes.max())
        else:
            self.true_lb = self.true_ub = 0
            align = 1
        if lb is None:
            lb = self.true_lb
        if extent is None:
            extent = self.true_ub - lb
            extent = (extent + align - 1) // align * align      # trailing padding, as a C struct would have
        self.lb, self.extent = int(lb), int(extent)
        self.combiner, self.contents = combiner, contents
        self.committed = False
        self.name = ""
        self.handle = -1
        self._blocks: Optional[Tuple[np.ndarray, np.ndarray]] = None

    def elem_sizes(self) -> np.ndarray:
        table = np.array([np.dtype(_NP[k]).itemsize for k in range(_NKINDS)], dtype=np.int64)
        return table[self.code] if self.code.size else np.zeros(0, np.int64)

    def blocks(self) -> Tuple[np.ndarray, np.ndarray]:
        """(displacement, length) of maximal contiguous byte runs, in type-map order."""
        if self._blocks is None:
            if self.disp.size == 0:
                self._blocks = (np.zeros(0, np.int64), np.zeros(0, np.int64))
            else:
                sz = self.elem_sizes()
                brk = np.ones(self.disp.size, dtype=bool)
                brk[1:] = self.disp[1:] != self.disp[:-1] + sz[:-1]
                starts = np.flatnonzero(brk)
                ends = np.append(starts[1:], self.disp.size)
                csum = np.concatenate([[0], np.cumsum(sz)])
                self._blocks = (np.ascontiguousarray(self.disp[starts]),
                                np.ascontiguousarray(csum[ends] - csum[starts]))
        return self._blocks


_TYPES: Dict[int, Datatype] = {}
for _k in range(10):
    _t = Datatype([0], [_k])
    _t.committed, _t.name, _t.handle = True, _BASIC_NAMES[_k], _k
    _TYPES[_k] = _t
_next_handle = [DERIVED_BASE + 100]        # DERIVED_BASE .. +99 are the predefined types below


def _predefine(handle: int, name: str, t: Datatype) -> int:
    t.committed, t.name, t.handle, t.combiner = True, name, handle, COMBINER_NAMED
    _TYPES[handle] = t
    return handle


def _pair(value_code: int, name: str, handle: int) -> int:
    """The {value, int} records MAXLOC / MINLOC reduce over, laid out like the C struct."""
    vsize = np.dtype(_NP[value_code]).itemsize
    ioff = (vsize + 3) // 4 * 4
    return _predefine(handle, name, Datatype([0, ioff], [value_code, 3]))


# Predefined types of /root/reference/lua-mpi.h:130-151 that the native layer has no wire id for.  They are handled
# like committed derived types (packed on the wire, reduced by the generic reducer in extras.reduce_packed).
UNSIGNED_SHORT = _predefine(DERIVED_BASE + 0, "MPI_UNSIGNED_SHORT", Datatype([0], [10]))
UNSIGNED_LONG = _predefine(DERIVED_BASE + 1, "MPI_UNSIGNED_LONG", Datatype([0], [11]))
LONG_DOUBLE = _predefine(DERIVED_BASE + 2, "MPI_LONG_DOUBLE", Datatype([0], [12]))
PACKED = _predefine(DERIVED_BASE + 3, "MPI_PACKED", Datatype([0], [0]))
FLOAT_INT = _pair(5, "MPI_FLOAT_INT", DERIVED_BASE + 4)
DOUBLE_INT = _pair(6, "MPI_DOUBLE_INT", DERIVED_BASE + 5)
LONG_INT = _pair(4, "MPI_LONG_INT", DERIVED_BASE + 6)
SHORT_INT = _pair(2, "MPI_SHORT_INT", DERIVED_BASE + 7)
LONG_DOUBLE_INT = _pair(12, "MPI_LONG_DOUBLE_INT", DERIVED_BASE + 8)
globals()["2INT"] = TWOINT = _pair(3, "MPI_2INT", DERIVED_BASE + 9)          # mpiT["2INT"] / mpiT.TWOINT
LONG_LONG_INT = 9                                                            # MPI's other name for LONG_LONG
# MPI-1 bound markers: zero-size members of Type_create_struct that pin the lower / upper bound
LB = _predefine(DERIVED_BASE + 10, "MPI_LB", Datatype([], []))
UB = _predefine(DERIVED_BASE + 11, "MPI_UB", Datatype([], []))


def packed_numpy_dtype(dtype: int, records: bool = False) -> np.dtype:
    """NumPy view of the packed stream of ``dtype``: the scalar dtype when every basic element is of one kind (the
    stream is then a plain array of them), a record ``(f0, f1, ...)`` per element otherwise."""
    t = get(dtype)
    if not records and t.code.size and np.all(t.code == t.code[0]):
        return np.dtype(_NP[int(t.code[0])])
    return np.dtype([(f"f{i}", _NP[int(c)]) for i, c in enumerate(t.code)])


def is_derived(dtype: int) -> bool:
    return isinstance(dtype, int) and dtype >= DERIVED_BASE


def get(dtype: int) -> Datatype:
    try:
        return _TYPES[dtype]
    except KeyError:
        raise _err(ERR_TYPE, f"datatype handle {dtype}") from None


def _register(t: Datatype) -> int:
    h = _next_handle[0]
    _next_handle[0] += 1
    t.handle = h
    _TYPES[h] = t
    return h


def _replicate(old: Datatype, offsets: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    offsets = np.asarray(offsets, dtype=np.int64).reshape(-1)
    disp = (offsets[:, None] + old.disp[None, :]).reshape(-1)
    code = np.tile(old.code, offsets.size)
    return disp, code


def _bounds(old: Datatype, offsets: np.ndarray) -> Tuple[int, int]:
    """lb / extent of a type made of copies of ``old`` at ``offsets`` (MPI: min lb, max ub over the copies)."""
    offsets = np.asarray(offsets, dtype=np.int64).reshape(-1)
    if offsets.size == 0:
        return 0, 0
    lb = int(offsets.min()) + old.lb
    ub = int(offsets.max()) + old.lb + old.extent
    return lb, ub - lb


# ---- constructors ------------------------------------------------------------------------------------------------
def Type_contiguous(count: int, oldtype: int) -> int:
    old = get(oldtype)
    off = np.arange(count, dtype=np.int64) * old.extent
    lb, ext = _bounds(old, off)
    return _register(Datatype(*_replicate(old, off), lb, ext, COMBINER_CONTIGUOUS, ([count], [], [oldtype])))


def _vector_offsets(count, blocklength, stride_bytes, old):
    return (np.arange(count, dtype=np.int64)[:, None] * stride_bytes +
            np.arange(blocklength, dtype=np.int64)[None, :] * old.extent).reshape(-1)


def Type_vector(count: int, blocklength: int, stride: int, oldtype: int) -> int:
    """``count`` blocks of ``blocklength`` elements, block starts ``stride`` ELEMENTS apart."""
    old = get(oldtype)
    off = _vector_offsets(count, blocklength, stride * old.extent, old)
    lb, ext = _bounds(old, off)
    return _register(Datatype(*_replicate(old, off), lb, ext, COMBINER_VECTOR,
                              ([count, blocklength, stride], [], [oldtype])))


def Type_create_hvector(count: int, blocklength: int, stride_bytes: int, oldtype: int) -> int:
    old = get(oldtype)
    off = _vector_offsets(count, blocklength, stride_bytes, old)
    lb, ext = _bounds(old, off)
    return _register(Datatype(*_replicate(old, off), lb, ext, COMBINER_HVECTOR,
                              ([count, blocklength], [stride_bytes], [oldtype])))


def _indexed_offsets(blocklengths, displs_bytes, old):
    parts = [d + np.arange(b, dtype=np.int64) * old.extent for b, d in zip(blocklengths, displs_bytes)]
    return np.concatenate(parts) if parts else np.zeros(0, np.int64)


def Type_indexed(blocklengths: Sequence[int], displacements: Sequence[int], oldtype: int) -> int:
    """Displacements in multiples of the old type's extent."""
    old = get(oldtype)
    off = _indexed_offsets(blocklengths, [d * old.extent for d in displacements], old)
    lb, ext = _bounds(old, off)
    return _register(Datatype(*_replicate(old, off), lb, ext, COMBINER_INDEXED,
                              ([len(blocklengths), *blocklengths, *displacements], [], [oldtype])))


def Type_create_hindexed(blocklengths: Sequence[int], displacements_bytes: Sequence[int], oldtype: int) -> int:
    old = get(oldtype)
    off = _indexed_offsets(blocklengths, list(displacements_bytes), old)
    lb, ext = _bounds(old, off)
    return _register(Datatype(*_replicate(old, off), lb, ext, COMBINER_HINDEXED,
                              ([len(blocklengths), *blocklengths], list(displacements_bytes), [oldtype])))


def Type_create_indexed_block(blocklength: int, displacements: Sequence[int], oldtype: int) -> int:
    old = get(oldtype)
    off = _indexed_offsets([blocklength] * len(displacements), [d * old.extent for d in displacements], old)
    lb, ext = _bounds(old, off)
    return _register(Datatype(*_replicate(old, off), lb, ext, COMBINER_INDEXED_BLOCK,
                              ([len(displacements), blocklength, *displacements], [], [oldtype])))


def Type_create_struct(blocklengths: Sequence[int], displacements_bytes: Sequence[int], types: Sequence[int]) -> int:
    disp, code, lbs, ubs = [], [], [], []
    forced_lb = forced_ub = None
    for b, d, t in zip(blocklengths, displacements_bytes, types):
        if t == LB:
            forced_lb = d if forced_lb is None else min(forced_lb, d)
            continue
        if t == UB:
            forced_ub = d if forced_ub is None else max(forced_ub, d)
            continue
        old = get(t)
        off = d + np.arange(b, dtype=np.int64) * old.extent
        dd, cc = _replicate(old, off)
        disp.append(dd)
        code.append(cc)
        if b:
            l, e = _bounds(old, off)
            lbs.append(l)
            ubs.append(l + e)
    disp = np.concatenate(disp) if disp else np.zeros(0, np.int64)
    code = np.concatenate(code) if code else np.zeros(0, np.int16)
    t = Datatype(disp, code, combiner=COMBINER_STRUCT,
                 contents=([len(blocklengths), *blocklengths], list(displacements_bytes), list(types)))
    if lbs:
        # bounds come from the members' bounds (a member may carry a resized extent); keep the alignment padding
        lb, ub = min(lbs), max(ubs)
        align = int(t.elem_sizes().max()) if t.disp.size else 1
        t.lb, t.extent = lb, ((ub - lb) + align - 1) // align * align
    if forced_lb is not None or forced_ub is not None:
        lb = t.lb if forced_lb is None else forced_lb
        ub = t.lb + t.extent if forced_ub is None else forced_ub
        t.lb, t.extent = lb, ub - lb
    return _register(t)


def Type_create_resized(oldtype: int, lb: int, extent: int) -> int:
    old = get(oldtype)
    return _register(Datatype(old.disp.copy(), old.code.copy(), lb, extent, COMBINER_RESIZED,
                              ([], [lb, extent], [oldtype])))


def Type_dup(oldtype: int) -> int:
    old = get(oldtype)
    t = Datatype(old.disp.copy(), old.code.copy(), old.lb, old.extent, COMBINER_DUP, ([], [], [oldtype]))
    t.committed, t.name = old.committed, old.name
    h = _register(t)
    from . import attrs
    attrs.copy_attrs("type", oldtype, h)
    return h


def Type_create_subarray(sizes: Sequence[int], subsizes: Sequence[int], starts: Sequence[int], order: int,
                         oldtype: int) -> int:
    """An n-dimensional sub-block of an n-dimensional array; the extent is the whole array."""
    old = get(oldtype)
    nd = len(sizes)
    if not (len(subsizes) == len(starts) == nd) or any(st < 0 or st + ss > sz for sz, ss, st in
                                                        zip(sizes, subsizes, starts)):
        raise _err(ERR_ARG, "Type_create_subarray")
    idx = [np.arange(st, st + ss, dtype=np.int64) for ss, st in zip(subsizes, starts)]
    off = _grid_offsets(sizes, idx, order) * old.extent
    total = int(np.prod(sizes)) * old.extent
    return _register(Datatype(*_replicate(old, off), old.lb, total, COMBINER_SUBARRAY,
                              ([nd, *sizes, *subsizes, *starts, order], [], [oldtype])))


def _grid_offsets(sizes: Sequence[int], idx: List[np.ndarray], order: int) -> np.ndarray:
    """Linear element offsets of the cartesian product of per-dimension index lists, in storage order."""
    nd = len(sizes)
    dims = list(range(nd)) if order == ORDER_C else list(range(nd - 1, -1, -1))   # slowest -> fastest
    strides = {}
    s = 1
    for d in reversed(dims):
        strides[d] = s
        s *= sizes[d]
    off = np.zeros(1, dtype=np.int64)
    for d in dims:
        off = (off[:, None] + idx[d][None, :] * strides[d]).reshape(-1)
    return off


def Type_create_darray(size: int, rank: int, gsizes: Sequence[int], distribs: Sequence[int], dargs: Sequence[int],
                       psizes: Sequence[int], order: int, oldtype: int) -> int:
    """The piece of a block / cyclic / block-cyclic distributed n-d array that process ``rank`` of a ``psizes``
    process grid owns (HPF-style, process grid always row-major as MPI specifies)."""
    old = get(oldtype)
    nd = len(gsizes)
    if int(np.prod(psizes)) != size:
        raise _err(ERR_ARG, "Type_create_darray: the process grid does not hold `size` processes")
    coords, r = [], rank
    for d in range(nd - 1, -1, -1):
        coords.insert(0, r % psizes[d])
        r //= psizes[d]
    idx = []
    for d in range(nd):
        g, p, c = gsizes[d], psizes[d], coords[d]
        if distribs[d] == DISTRIBUTE_NONE:
            if p != 1:
                raise _err(ERR_ARG, "Type_create_darray: DISTRIBUTE_NONE needs a process-grid extent of 1")
            idx.append(np.arange(g, dtype=np.int64))
            continue
        if distribs[d] == DISTRIBUTE_BLOCK:
            b = (g + p - 1) // p if dargs[d] == DISTRIBUTE_DFLT_DARG else dargs[d]
            if b * p < g:
                raise _err(ERR_ARG, "Type_create_darray: block size too small")
            idx.append(np.arange(min(c * b, g), min((c + 1) * b, g), dtype=np.int64))
        elif distribs[d] == DISTRIBUTE_CYCLIC:
            b = 1 if dargs[d] == DISTRIBUTE_DFLT_DARG else dargs[d]
            allidx = np.arange(g, dtype=np.int64)
            idx.append(allidx[(allidx // b) % p == c])
        else:
            raise _err(ERR_ARG, "Type_create_darray: unknown distribution")
    off = _grid_offsets(gsizes, idx, order) * old.extent
    total = int(np.prod(gsizes)) * old.extent
    return _register(Datatype(*_replicate(old, off), old.lb, total, COMBINER_DARRAY,
                              ([size, rank, nd, *gsizes, *distribs, *dargs, *psizes, order], [], [oldtype])))


def Type_commit(dtype: int) -> int:
    t = get(dtype)
    t.blocks()
    t.committed = True
    return dtype


def Type_free(dtype: int) -> int:
    """Releases a derived type; returns DATATYPE_NULL (-1).  Basic types cannot be freed."""
    if not is_derived(dtype) or dtype < DERIVED_BASE + 100:
        raise _err(ERR_TYPE, "Type_free of a predefined datatype")
    from . import attrs
    attrs.drop_object("type", dtype)
    _TYPES.pop(dtype, None)
    return -1


# ---- queries -----------------------------------------------------------------------------------------------------
def Type_size(dtype: int) -> int:
    return get(dtype).size


def Type_get_extent(dtype: int) -> Tuple[int, int]:
    t = get(dtype)
    return t.lb, t.extent


def Type_get_true_extent(dtype: int) -> Tuple[int, int]:
    t = get(dtype)
    return t.true_lb, t.true_ub - t.true_lb


def Type_get_envelope(dtype: int) -> Tuple[int, int, int, int]:
    """-> (num_integers, num_addresses, num_datatypes, combiner)."""
    t = get(dtype)
    i, a, d = t.contents
    return len(i), len(a), len(d), t.combiner


def Type_get_contents(dtype: int) -> Tuple[List[int], List[int], List[int]]:
    t = get(dtype)
    if t.combiner == COMBINER_NAMED:
        raise _err(ERR_TYPE, "Type_get_contents of a predefined datatype")
    i, a, d = t.contents
    return list(i), list(a), list(d)


def Type_set_name(dtype: int, name: str) -> None:
    get(dtype).name = name[:128]


def Type_get_name(dtype: int) -> str:
    return get(dtype).name


def Type_match_size(typeclass: int, size: int) -> int:
    """The predefined type of a class with the given size in bytes."""
    table = {TYPECLASS_REAL: {4: 5, 8: 6}, TYPECLASS_INTEGER: {1: 1, 2: 2, 4: 3, 8: 4}}
    try:
        return table[typeclass][size]
    except KeyError:
        raise _err(ERR_ARG, "Type_match_size") from None


def Get_address(location) -> int:
    """Address of a tensor / array's first element (an MPI_Aint).  Differences of addresses are the byte
    displacements that ``Type_create_struct`` / ``Type_create_hindexed`` take."""
    if isinstance(location, torch.Tensor):
        return location.data_ptr()
    if isinstance(location, np.ndarray):
        return location.ctypes.data
    raise TypeError(f"Get_address of {type(location)}")


def Get_elements(status, dtype: int) -> int:
    """Number of BASIC elements a receive delivered (can end inside a derived element)."""
    t = get(dtype)
    nbytes = int(status.count_bytes)
    if t.size == 0:
        return 0
    full, rest = divmod(nbytes, t.size)
    csum = np.cumsum(t.elem_sizes())
    part = int(np.searchsorted(csum, rest, side="right"))
    if rest and (part == 0 or csum[part - 1] != rest):
        return -32766                                           # UNDEFINED: ends inside a basic element
    return full * t.code.size + part


def count_from_bytes(nbytes: int, dtype: int) -> int:
    """Get_count for any datatype: whole elements, or UNDEFINED."""
    t = get(dtype)
    if t.size == 0:
        return 0
    return nbytes // t.size if nbytes % t.size == 0 else -32766


# ---- data movement -----------------------------------------------------------------------------------------------
def _host_bytes(buf, writable: bool) -> np.ndarray:
    """uint8 numpy view aliasing a host buffer's memory."""
    if isinstance(buf, torch.Tensor):
        if not buf.is_contiguous():
            raise ValueError("buffers must be contiguous (describe strides with a datatype instead)")
        return buf.view(-1).view(torch.uint8).numpy()
    if isinstance(buf, np.ndarray):
        if not buf.flags["C_CONTIGUOUS"]:
            raise ValueError("buffers must be contiguous (describe strides with a datatype instead)")
        return buf.reshape(-1).view(np.uint8)
    if isinstance(buf, (bytearray, memoryview)):
        return np.frombuffer(buf, dtype=np.uint8)
    if isinstance(buf, bytes) and not writable:
        return np.frombuffer(buf, dtype=np.uint8)
    raise TypeError(f"unsupported buffer type {type(buf)}")


def _span(t: Datatype, count: int) -> Tuple[int, int]:
    """[lo, hi) byte range touched by ``count`` instances."""
    if count == 0 or t.disp.size == 0:
        return 0, 0
    ext_lo, ext_hi = min(0, (count - 1) * t.extent), max(0, (count - 1) * t.extent)
    return t.true_lb + ext_lo, t.true_ub + ext_hi


def _check(t: Datatype, count: int, have_bytes: int, what: str):
    if not t.committed:
        raise _err(ERR_TYPE, f"{what}: datatype is not committed")
    lo, hi = _span(t, count)
    if lo < 0:
        raise _err(ERR_ARG, f"{what}: the type map reaches {-lo} bytes before the buffer address")
    if hi > have_bytes:
        raise _err(ERR_TRUNCATE, f"{what}: {count} x datatype spans {hi} bytes, the buffer holds {have_bytes}")


def _cptr(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def pack_bytes(buf, count: int, dtype: int) -> np.ndarray:
    """Packed stream (new uint8 array) of ``count`` elements of ``dtype`` read from ``buf``."""
    t = get(dtype)
    nbytes = count * t.size
    if isinstance(buf, torch.Tensor) and buf.is_cuda:
        flat = buf.view(-1).view(torch.uint8)
        _check(t, count, flat.numel(), "Pack")
        out = torch.empty(nbytes, dtype=torch.uint8, device=buf.device)
        _device_move(t, flat, out, nbytes, to_stream=True)
        return out.cpu().numpy()
    src = _host_bytes(buf, False)
    _check(t, count, src.size, "Pack")
    out = np.empty(nbytes, dtype=np.uint8)
    d, n = t.blocks()
    host().mpit_dt_pack(_cptr(src), count, t.extent, d.size, _cptr(d), _cptr(n), _cptr(out), nbytes)
    return out


def unpack_bytes(stream: np.ndarray, buf, count: int, dtype: int, limit: Optional[int] = None) -> int:
    """Scatter a packed stream into ``buf``; returns the bytes consumed."""
    t = get(dtype)
    nbytes = count * t.size if limit is None else min(limit, count * t.size)
    stream = np.ascontiguousarray(stream[:nbytes])
    if isinstance(buf, torch.Tensor) and buf.is_cuda:
        flat = buf.view(-1).view(torch.uint8)
        _check(t, count, flat.numel(), "Unpack")
        staged = torch.from_numpy(stream).to(buf.device)
        _device_move(t, flat, staged, nbytes, to_stream=False)
        return nbytes
    dst = _host_bytes(buf, True)
    _check(t, count, dst.size, "Unpack")
    d, n = t.blocks()
    return int(host().mpit_dt_unpack(_cptr(dst), count, t.extent, d.size, _cptr(d), _cptr(n), _cptr(stream), nbytes))


def _device_move(t: Datatype, strided: torch.Tensor, stream: torch.Tensor, nbytes: int, to_stream: bool) -> None:
    """Gather / scatter on the GPU (``mpit_dt_move_device``, csrc/kernels/dtpack.cu).  The block tables are uploaded
    once per datatype and device; the word size is the largest power of two that every address involved is a
    multiple of."""
    import math
    from .._native import kernels, check, ptr, cur_stream
    if nbytes == 0:
        return
    dev = strided.device
    if not hasattr(t, "_dev_tables"):
        t._dev_tables = {}
    if dev not in t._dev_tables:
        d, n = t.blocks()
        pre = np.concatenate([[0], np.cumsum(n)]).astype(np.int64)
        unit = 16
        for v in (*d.tolist(), *n.tolist(), t.extent):
            unit = math.gcd(unit, int(v))
        regular = None
        if d.size and np.all(n == n[0]) and (d.size == 1 or np.all(np.diff(d) == d[1] - d[0])):
            regular = (int(d[0]), int(n[0]), int(d[1] - d[0]) if d.size > 1 else int(n[0]))
        t._dev_tables[dev] = (torch.from_numpy(d.copy()).to(dev), torch.from_numpy(pre).to(dev), unit, int(d.size),
                              regular)
    d_disp, d_pre, unit, nblk, regular = t._dev_tables[dev]
    unit = math.gcd(math.gcd(unit, nbytes), math.gcd(strided.data_ptr() % 16 or 16, stream.data_ptr() % 16 or 16))
    with torch.cuda.device(dev):
        if regular is not None:
            d0, blen, bstride = regular
            check(kernels().mpit_dt_move_device_regular(ptr(strided), ptr(stream), nbytes, t.size, t.extent, d0, blen,
                                                        bstride, unit, int(to_stream), cur_stream()),
                  "mpit_dt_move_device_regular")
        else:
            check(kernels().mpit_dt_move_device(ptr(strided), ptr(stream), nbytes, t.size, t.extent, nblk,
                                                ptr(d_disp), ptr(d_pre), unit, int(to_stream), cur_stream()),
                  "mpit_dt_move_device")


def Pack_size(incount: int, dtype: int, comm: int = 0) -> int:
    return incount * get(dtype).size


def Pack(inbuf, incount: int, dtype: int, outbuf, position: int, comm: int = 0) -> int:
    """Append ``incount`` elements to the packed buffer ``outbuf`` at byte ``position``; returns the new position."""
    stream = pack_bytes(inbuf, incount, dtype)
    out = _host_bytes(outbuf, True)
    if position + stream.size > out.size:
        raise _err(ERR_TRUNCATE, "Pack: the output buffer is too small (see Pack_size)")
    out[position:position + stream.size] = stream
    return position + stream.size


def Unpack(inbuf, position: int, outbuf, outcount: int, dtype: int, comm: int = 0) -> int:
    src = _host_bytes(inbuf, False)
    need = outcount * get(dtype).size
    if position + need > src.size:
        raise _err(ERR_TRUNCATE, "Unpack: the packed buffer ends before the requested elements")
    unpack_bytes(src[position:position + need], outbuf, outcount, dtype)
    return position + need


# ---- external32: a portable representation (big endian, fixed sizes) ---------------------------------------------
def _ext_layout(t: Datatype, count: int):
    code = np.tile(t.code, count)
    nat = t.elem_sizes()
    nat = np.tile(nat, count)
    ext = np.array([np.dtype(_EXT32[k]).itemsize for k in range(_NKINDS)], dtype=np.int64)[code]
    return code, nat, np.cumsum(nat) - nat, ext, np.cumsum(ext) - ext


def _check_rep(datarep: str):
    if datarep != "external32":
        raise _err(ERR_UNSUPPORTED_DATAREP, f"data representation {datarep!r}")


def Pack_external_size(datarep: str, incount: int, dtype: int) -> int:
    _check_rep(datarep)
    t = get(dtype)
    ext = np.array([np.dtype(_EXT32[k]).itemsize for k in range(_NKINDS)], dtype=np.int64)
    return incount * int(ext[t.code].sum())


def Pack_external(datarep: str, inbuf, incount: int, dtype: int, outbuf, position: int) -> int:
    _check_rep(datarep)
    t = get(dtype)
    native = pack_bytes(inbuf, incount, dtype)
    code, nat, nat_off, ext, ext_off = _ext_layout(t, incount)
    out = _host_bytes(outbuf, True)
    total = int(ext.sum())
    if position + total > out.size:
        raise _err(ERR_TRUNCATE, "Pack_external: the output buffer is too small (see Pack_external_size)")
    dst = out[position:position + total]
    for k in np.unique(code):
        sel = code == k
        ns, es = int(nat[sel][0]), int(ext[sel][0])
        vals = native[(nat_off[sel][:, None] + np.arange(ns)[None, :]).reshape(-1)].view(_NP[int(k)])
        dst[(ext_off[sel][:, None] + np.arange(es)[None, :]).reshape(-1)] = vals.astype(_EXT32[int(k)]).view(np.uint8)
    return position + total


def Unpack_external(datarep: str, inbuf, position: int, outbuf, outcount: int, dtype: int) -> int:
    _check_rep(datarep)
    t = get(dtype)
    code, nat, nat_off, ext, ext_off = _ext_layout(t, outcount)
    src = _host_bytes(inbuf, False)
    total = int(ext.sum())
    if position + total > src.size:
        raise _err(ERR_TRUNCATE, "Unpack_external: the packed buffer ends before the requested elements")
    src = src[position:position + total]
    native = np.empty(int(nat.sum()), dtype=np.uint8)
    for k in np.unique(code):
        sel = code == k
        ns, es = int(nat[sel][0]), int(ext[sel][0])
        vals = np.ascontiguousarray(src[(ext_off[sel][:, None] + np.arange(es)[None, :]).reshape(-1)]).view(_EXT32[int(k)])
        native[(nat_off[sel][:, None] + np.arange(ns)[None, :]).reshape(-1)] = vals.astype(_NP[int(k)]).view(np.uint8)
    unpack_bytes(native, outbuf, outcount, dtype)
    return position + total


# ---- hook used by the message layer's buffer resolution (mpit_b200.msg._Buf) -------------------------------------
def stage_for_message(buf, count: int, dtype: int, writable: bool):
    """-> (uint8 staging array, completion hook).  A send travels as the packed stream; a receive lands in a staging
    stream that was pre-filled from the user buffer (so bytes a short message does not deliver are written back
    unchanged) and is scattered by the hook once the transfer has completed."""
    stream = pack_bytes(buf, count, dtype)
    if not writable:
        return stream, None
    return stream, (lambda: unpack_bytes(stream, buf, count, dtype))


__all__ = [_n for _n in dir() if _n[0].isupper() and _n not in ("Any", "Callable", "Dict", "List", "Optional", "Sequence",
                                                                 "Tuple")] + ["2INT"]
