#!/usr/bin/env python
"""In-tree native build for mpit_b200 (sm_100a only).

Two shared libraries are produced under ``mpit_b200/lib``:

* ``libmpit_kernels.so`` -- every CUDA kernel (C ABI, see ``include/mpit_b200.h``)
  plus the symmetric-heap runtime, compiled with
  ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo``.
* ``libmpit_host.so``    -- the C++ host message layer (``include/mpit_msg.h``),
  no CUDA dependency, so the CPU control plane runs on boxes without a GPU.

The reference builds one C translation unit with CMake + luarocks against an MPI
install (/root/reference/CMakeLists.txt:1-29); here the build is a single Python
script so that ``__graft_entry__.build()`` can drive it and the resulting ``.so``
files travel with the tree to the GPU box.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INC = ROOT / "include"
OUT = ROOT / "mpit_b200" / "lib"
OBJ = ROOT / "build" / "obj"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function",
    "-Xptxas", "-v",
    f"-I{INC}", f"-I{CSRC}",
]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-pthread", f"-I{INC}", f"-I{CSRC}"]


def _digest(path: Path, extra: str) -> str:
    h = hashlib.sha256()
    h.update(extra.encode())
    h.update(path.read_bytes())
    for hdr in sorted(list(INC.glob("*.h")) + list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.h"))):
        h.update(hdr.read_bytes())
    return h.hexdigest()


def _compile(src: Path, verbose: bool) -> Path:
    is_cu = src.suffix == ".cu"
    cmd_head = [NVCC, *NVCC_FLAGS] if is_cu else [CXX, *CXX_FLAGS]
    obj = OBJ / (src.relative_to(CSRC).as_posix().replace("/", "__") + ".o")
    stamp = obj.with_suffix(".sha")
    dig = _digest(src, " ".join(cmd_head))
    if obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return obj
    cmd = [*cmd_head, "-c", str(src), "-o", str(obj)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = obj.with_suffix(".log")
    log.write_text(res.stdout + res.stderr)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(f"compile failed: {src}")
    if verbose:
        sys.stderr.write(f"[build] {src.relative_to(ROOT)} ok\n")
    stamp.write_text(dig)
    return obj


def build(verbose: bool = True, jobs: int | None = None) -> dict:
    OUT.mkdir(parents=True, exist_ok=True)
    OBJ.mkdir(parents=True, exist_ok=True)
    if shutil.which(NVCC) is None and not Path(NVCC).exists():
        raise RuntimeError("nvcc not found")
    kern_srcs = sorted((CSRC / "kernels").glob("*.cu")) + sorted((CSRC / "symm").glob("*.cu"))
    host_srcs = sorted((CSRC / "msg").glob("*.cpp")) + sorted((CSRC / "io").glob("*.cpp"))
    jobs = jobs or min(8, os.cpu_count() or 4)
    with ThreadPoolExecutor(jobs) as ex:
        kern_objs = list(ex.map(lambda s: _compile(s, verbose), kern_srcs))
        host_objs = list(ex.map(lambda s: _compile(s, verbose), host_srcs))
    libs = {}
    if kern_objs:
        lib = OUT / "libmpit_kernels.so"
        cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a",
               "-o", str(lib), *map(str, kern_objs), "-lpthread", "-ldl", "-lrt"]
        subprocess.run(cmd, check=True)
        libs["kernels"] = lib
    if host_objs:
        lib = OUT / "libmpit_host.so"
        cmd = [CXX, "-shared", "-o", str(lib), *map(str, host_objs), "-pthread", "-lrt"]
        subprocess.run(cmd, check=True)
        libs["host"] = lib
    if verbose:
        for k, v in libs.items():
            sys.stderr.write(f"[build] {k}: {v.relative_to(ROOT)} ({v.stat().st_size >> 10} KiB)\n")
    return libs


def ptxas_report() -> str:
    """Concatenate the per-file ``-Xptxas -v`` logs (registers / spills / smem)."""
    out = []
    for log in sorted(OBJ.glob("*.log")):
        out.append(f"==== {log.stem} ====\n{log.read_text()}")
    return "\n".join(out)


if __name__ == "__main__":
    build(verbose=True)
    if "--ptxas" in sys.argv:
        print(ptxas_report())
