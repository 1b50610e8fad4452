"""Bind a C header's ``int f(...)`` declarations onto a ctypes library.

The reference generates its 258 Lua->C wrappers from JSON specs with ``readspec.py``
(/root/reference/readspec.py:73-188, type map :105-124).  The equivalent step here parses the
C ABI headers in ``include/`` and sets ``argtypes``/``restype`` on the loaded library, so the
header is the single source of truth for the binding.
"""
from __future__ import annotations

import ctypes
import re
from pathlib import Path

_INCLUDE = Path(__file__).resolve().parent.parent / "include"

_SCALARS = {
    "int": ctypes.c_int,
    "unsigned": ctypes.c_uint,
    "unsigned int": ctypes.c_uint,
    "long long": ctypes.c_longlong,
    "unsigned long long": ctypes.c_ulonglong,
    "long": ctypes.c_long,
    "size_t": ctypes.c_size_t,
    "float": ctypes.c_float,
    "double": ctypes.c_double,
    "int64_t": ctypes.c_int64,
    "uint64_t": ctypes.c_uint64,
    "int32_t": ctypes.c_int32,
    "uint32_t": ctypes.c_uint32,
}

_DECL = re.compile(r"^\s*(int|double|void|long long|const char\*|void\*)\s+(mpit_\w+)\s*\(([^;{]*)\)\s*;", re.M | re.S)


def _ctype(arg: str):
    arg = arg.strip()
    if arg in ("void", ""):
        return None
    if "*" in arg:
        base = arg.split("*")[0].replace("const", "").strip()
        if base == "char" and arg.count("*") == 1:
            return ctypes.c_char_p
        return ctypes.c_void_p
    # strip the parameter name
    toks = arg.replace("const", "").split()
    ty = " ".join(toks[:-1]) if len(toks) > 1 else toks[0]
    if ty not in _SCALARS:
        raise ValueError(f"unknown C type in header: {arg!r}")
    return _SCALARS[ty]


def parse_header(name: str) -> dict:
    text = (_INCLUDE / name).read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    out = {}
    for m in _DECL.finditer(text):
        ret, fn, args = m.group(1), m.group(2), m.group(3)
        argtypes = [t for t in (_ctype(a) for a in args.split(",")) if t is not None]
        restype = {"int": ctypes.c_int, "double": ctypes.c_double, "void": None,
                   "long long": ctypes.c_longlong, "const char*": ctypes.c_char_p,
                   "void*": ctypes.c_void_p}[ret]
        out[fn] = (restype, argtypes)
    return out


def bind(lib: ctypes.CDLL, header: str) -> list:
    """Apply the header's prototypes to ``lib``; returns the list of bound names."""
    names = []
    for fn, (restype, argtypes) in parse_header(header).items():
        try:
            f = getattr(lib, fn)
        except AttributeError as e:  # declared but not built: fail loudly
            raise RuntimeError(f"{header} declares {fn} but the library does not export it") from e
        f.restype = restype
        f.argtypes = argtypes
        names.append(fn)
    return names
