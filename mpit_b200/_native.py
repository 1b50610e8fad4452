"""ctypes loader for the in-tree native libraries.

``kernels()`` returns ``libmpit_kernels.so`` (CUDA, C ABI of ``include/mpit_b200.h``) and
``host()`` returns ``libmpit_host.so`` (the C++ message layer, ``include/mpit_msg.h``).
Both are built by ``tools/build.py`` and live in ``mpit_b200/lib`` so that they travel with
the source tree.  There is deliberately no Python/PyTorch fallback for a missing kernel
library on a GPU box: ops raise instead of silently running eager code.

The reference exposes its C layer to Lua through ``luaL_register`` tables
(/root/reference/lua-mpi.h:235-297); ctypes over a flat C ABI plays that role here.
"""
from __future__ import annotations

import ctypes
import os
from functools import lru_cache
from pathlib import Path

_LIBDIR = Path(__file__).resolve().parent / "lib"


class NativeError(RuntimeError):
    pass


def _load(name: str) -> ctypes.CDLL:
    path = _LIBDIR / name
    if not path.exists():
        if os.environ.get("MPIT_AUTOBUILD", "1") == "1":
            import importlib.util
            spec = importlib.util.spec_from_file_location(
                "mpit_build", Path(__file__).resolve().parent.parent / "tools" / "build.py")
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            mod.build(verbose=False)
        if not path.exists():
            raise NativeError(f"{path} is missing: run `python tools/build.py`")
    return ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)


@lru_cache(maxsize=None)
def kernels() -> ctypes.CDLL:
    from ._cabi import bind
    lib = _load("libmpit_kernels.so")
    bind(lib, "mpit_b200.h")
    if os.environ.get("MPIT_PDL", "1") == "0":     # programmatic dependent launch off (A/B runs, debugging)
        lib.mpit_set_pdl(0)
    return lib


@lru_cache(maxsize=None)
def host() -> ctypes.CDLL:
    from ._cabi import bind
    lib = _load("libmpit_host.so")
    bind(lib, "mpit_msg.h")
    return lib


class _LaunchCounter:
    """Counts launches of our own kernels (every successful native kernel call is >= 1 launch).

    ``bench.py`` reports this as ``gpu_launches``; graph replays add the number of launches
    that were captured into the graph.
    """
    count = 0


launches = _LaunchCounter()


def check(code: int, what: str = "") -> None:
    """Raise on a non-zero return code from a native entry point."""
    if code == 0:
        launches.count += 1
        return
    msg = {-2: "unsupported", -3: "timeout", -4: "bad argument",
           -5: "request exceeds the built-in scratch of a no-allocation entry point"}.get(code)
    if msg is None and code > 0:
        try:
            import torch
            msg = f"CUDA error {code}"
            # cudaGetErrorString through torch's cudart handle
            msg = f"CUDA error {code}: {torch.cuda.cudart().cudaGetErrorString(code)}"  # type: ignore[attr-defined]
        except Exception:
            pass
    raise NativeError(f"{what or 'native call'} failed: {msg or code}")


def ptr(t) -> ctypes.c_void_p:
    """Device/host pointer of a tensor (None -> NULL)."""
    if t is None:
        return ctypes.c_void_p(0)
    return ctypes.c_void_p(t.data_ptr())


def cur_stream() -> ctypes.c_void_p:
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


c_f = ctypes.c_float
c_i = ctypes.c_int
c_ll = ctypes.c_longlong
c_ull = ctypes.c_ulonglong
