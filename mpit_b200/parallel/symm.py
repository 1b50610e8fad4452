"""Symmetric heap front-end (csrc/symm/symm_heap.cu).

Every rank allocates a buffer of the same size; each buffer is mapped into every other rank's
address space (CUDA IPC, or VMM + POSIX fds when NVLS multicast is wanted), so kernels can
``ld``/``st``/``red`` straight into a peer's HBM over NVLink.  Handle exchange -- the only
collective step -- goes through ``torch.distributed`` (any backend; object collectives), i.e. the
control plane; nothing on the hot path calls it.

Reference counterpart: CUDA-aware MPI given a raw device pointer (/root/reference/lua-mpi.h:70-78).
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass, field
from typing import List, Optional

import torch

from .._native import kernels, check, NativeError

MAX_RANKS = 8


class _RawCuda:
    """Adapter exposing a raw device pointer through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2, "strides": None}


def tensor_from_ptr(ptr: int, nbytes: int, device: torch.device, dtype=torch.float32) -> torch.Tensor:
    t = torch.as_tensor(_RawCuda(ptr, nbytes), device=device)
    return t.view(dtype)


@dataclass
class SymmBuffer:
    local: torch.Tensor              # this rank's copy (zero-initialised)
    ptrs: List[int]                  # device address of every rank's copy, as mapped here
    rank: int
    nranks: int
    mc_ptr: int = 0                  # NVLS multicast address (0 = none)
    _keep: list = field(default_factory=list, repr=False)

    def ptr_array(self):
        arr = (ctypes.c_ulonglong * MAX_RANKS)()
        for i, p in enumerate(self.ptrs):
            arr[i] = p
        return arr

    def offset_ptr_array(self, byte_offset: int):
        arr = (ctypes.c_ulonglong * MAX_RANKS)()
        for i, p in enumerate(self.ptrs):
            arr[i] = p + byte_offset
        return arr


def _mpit_world():
    """The host message layer as control plane (jobs started with ``python -m mpit_b200.launch.run``, where
    torch.distributed is not initialised): (rank, size) or None."""
    try:
        from .. import msg as mpiT
        if mpiT.Initialized() and mpiT.get_size() > 1:
            return mpiT.get_rank(), mpiT.get_size()
    except Exception:
        pass
    return None


def _dist_info(group=None):
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    w = _mpit_world()
    return w if w is not None else (0, 1)


def _all_gather_obj(obj, group=None):
    import torch.distributed as dist
    rank, world = _dist_info(group)
    if world == 1:
        return [obj]
    if dist.is_available() and dist.is_initialized():
        out = [None] * world
        dist.all_gather_object(out, obj, group=group)
        return out
    # host message layer: lengths with Allgather, pickles with a padded Allgather
    import pickle
    from .. import msg as mpiT
    blob = pickle.dumps(obj)
    lens = torch.zeros(world, dtype=torch.int64)
    mpiT.Allgather(torch.tensor([len(blob)], dtype=torch.int64), 1, mpiT.LONG, lens, mpiT.COMM_WORLD)
    width = int(lens.max())
    mine = torch.zeros(width, dtype=torch.uint8)
    mine[:len(blob)] = torch.frombuffer(bytearray(blob), dtype=torch.uint8)
    allb = torch.zeros(width * world, dtype=torch.uint8)
    mpiT.Allgather(mine, width, mpiT.BYTE, allb, mpiT.COMM_WORLD)
    return [pickle.loads(bytes(allb[r * width:r * width + int(lens[r])].tolist())) for r in range(world)]


def _barrier(group=None):
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        if dist.get_world_size(group) > 1:
            dist.barrier(group=group)
        return
    if _mpit_world() is not None:
        from .. import msg as mpiT
        mpiT.Barrier(mpiT.COMM_WORLD)


class SymmetricHeap:
    """One-process-per-GPU symmetric allocator.

    ``mode='ipc'`` shares cudaMalloc memory through CUDA IPC handles; ``mode='vmm'`` uses
    cuMemCreate + fd export and, when the device supports it, also builds an NVLS multicast
    mapping (``SymmBuffer.mc_ptr``).  ``mode='auto'`` tries vmm and falls back to ipc.
    """

    def __init__(self, device: Optional[torch.device] = None, group=None, mode: str = "auto"):
        self.k = kernels()
        self.group = group
        self.rank, self.nranks = _dist_info(group)
        if self.nranks > MAX_RANKS:
            raise ValueError(f"at most {MAX_RANKS} ranks per box")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        torch.cuda.set_device(self.dev_index)
        torch.cuda.init()
        mode = os.environ.get("MPIT_SYMM_MODE", mode)
        if mode == "auto":
            g = self.k.mpit_vmm_granularity(self.dev_index, self.nranks, 0)
            mode = "vmm" if g > 0 else "ipc"
            modes = _all_gather_obj(mode, group)
            mode = "vmm" if all(m == "vmm" for m in modes) else "ipc"
        self.mode = mode
        self.multicast_ok = False
        if mode == "vmm" and self.nranks > 1 and os.environ.get("MPIT_NO_MULTICAST", "0") != "1":
            ok = bool(self.k.mpit_vmm_multicast_supported(self.dev_index))
            self.multicast_ok = all(_all_gather_obj(ok, group))
        self._signal: Optional[SymmBuffer] = None

    # -- allocation ----------------------------------------------------------------------------
    def alloc(self, numel: int, dtype=torch.float32, multicast: bool = False) -> SymmBuffer:
        esz = torch.empty((), dtype=dtype).element_size()
        nbytes = max(256, (numel * esz + 255) // 256 * 256)
        if self.mode == "vmm":
            return self._alloc_vmm(nbytes, dtype, multicast and self.multicast_ok)
        return self._alloc_ipc(nbytes, dtype)

    def _alloc_ipc(self, nbytes: int, dtype) -> SymmBuffer:
        k = self.k
        p = ctypes.c_ulonglong(0)
        check(k.mpit_symm_malloc(nbytes, ctypes.byref(p)), "mpit_symm_malloc")
        handle = (ctypes.c_char * 64)()
        if self.nranks > 1:
            check(k.mpit_symm_ipc_export(p.value, handle), "mpit_symm_ipc_export")
        handles = _all_gather_obj(bytes(handle), self.group)
        ptrs = []
        for r, h in enumerate(handles):
            if r == self.rank:
                ptrs.append(p.value)
            else:
                q = ctypes.c_ulonglong(0)
                buf = (ctypes.c_char * 64).from_buffer_copy(h)
                check(k.mpit_symm_ipc_import(buf, ctypes.byref(q)), "mpit_symm_ipc_import")
                ptrs.append(q.value)
        local = tensor_from_ptr(p.value, nbytes, self.device, dtype)
        _barrier(self.group)
        return SymmBuffer(local, ptrs, self.rank, self.nranks)

    def _alloc_vmm(self, nbytes: int, dtype, multicast: bool) -> SymmBuffer:
        k = self.k
        gran = k.mpit_vmm_granularity(self.dev_index, self.nranks, int(multicast))
        if gran <= 0:
            raise NativeError("VMM allocation is not available on this device")
        size = (nbytes + gran - 1) // gran * gran
        p = ctypes.c_ulonglong(0)
        fd = ctypes.c_int(-1)
        check(k.mpit_vmm_alloc(self.dev_index, size, ctypes.byref(p), ctypes.byref(fd)), "mpit_vmm_alloc")
        infos = _all_gather_obj((os.getpid(), fd.value), self.group)
        ptrs = []
        for r, (pid, rfd) in enumerate(infos):
            if r == self.rank:
                ptrs.append(p.value)
                continue
            lfd = k.mpit_fd_from_pid(pid, rfd)
            if lfd < 0:
                raise NativeError("pidfd_getfd failed: cannot import a peer's allocation")
            q = ctypes.c_ulonglong(0)
            check(k.mpit_vmm_import(self.dev_index, lfd, size, ctypes.byref(q)), "mpit_vmm_import")
            k.mpit_close_fd(lfd)
            ptrs.append(q.value)
        mc_ptr = 0
        if multicast and self.nranks > 1:
            mc_ptr = self._bind_multicast(p.value, size)
        _barrier(self.group)  # peers keep their fds open until everyone has imported
        local = tensor_from_ptr(p.value, nbytes, self.device, dtype)
        return SymmBuffer(local, ptrs, self.rank, self.nranks, mc_ptr)

    def _bind_multicast(self, local_ptr: int, size: int) -> int:
        k = self.k
        h = ctypes.c_ulonglong(0)
        fd = ctypes.c_int(-1)
        if self.rank == 0:
            check(k.mpit_mc_create(self.nranks, size, ctypes.byref(h), ctypes.byref(fd)), "mpit_mc_create")
        info = _all_gather_obj((os.getpid(), fd.value), self.group)[0]
        if self.rank != 0:
            lfd = k.mpit_fd_from_pid(info[0], info[1])
            if lfd < 0:
                raise NativeError("pidfd_getfd failed for the multicast object")
            check(k.mpit_mc_import(lfd, ctypes.byref(h)), "mpit_mc_import")
            k.mpit_close_fd(lfd)
        check(k.mpit_mc_add_device(h.value, self.dev_index), "mpit_mc_add_device")
        _barrier(self.group)   # every device added before any bind
        check(k.mpit_mc_bind(h.value, local_ptr, size), "mpit_mc_bind")
        _barrier(self.group)
        mc = ctypes.c_ulonglong(0)
        check(k.mpit_mc_map(h.value, self.dev_index, size, ctypes.byref(mc)), "mpit_mc_map")
        return mc.value

    # -- signal pads ----------------------------------------------------------------------------
    def signal_pad(self) -> SymmBuffer:
        if self._signal is None:
            self._signal = self.alloc(self.k.mpit_signal_words(), dtype=torch.int32)
        return self._signal

    def check_timeout(self) -> None:
        """Raise if a bounded device-side spin timed out (a peer never arrived)."""
        sig = self.signal_pad()
        err = int(sig.local[self.k.mpit_signal_words() - 64].item())
        if err != 0:
            raise NativeError("device barrier timed out: a peer rank did not launch its kernel")


class LocalPeerHeap:
    """Single-process stand-in: one 'rank' per visible GPU, peer access enabled between them.

    Lets one pytest process exercise the exchange kernels on 2+ GPUs (or on 1 GPU with every
    rank aliased to the same device, which is how the single-GPU CI tier covers the
    multi-shard arithmetic; barrier kernels are only used when every rank has its own GPU).
    """

    def __init__(self, nranks: int, devices: Optional[List[int]] = None):
        self.k = kernels()
        ndev = torch.cuda.device_count()
        self.devices = devices or [r % ndev for r in range(nranks)]
        self.nranks = nranks
        for a in set(self.devices):
            for b in set(self.devices):
                if a != b:
                    check(self.k.mpit_enable_peer_access(a, b), "mpit_enable_peer_access")

    def alloc(self, numel: int, dtype=torch.float32) -> List[SymmBuffer]:
        locs = [torch.zeros(max(numel, 64), dtype=dtype, device=f"cuda:{d}") for d in self.devices]
        ptrs = [t.data_ptr() for t in locs]
        return [SymmBuffer(locs[r], list(ptrs), r, self.nranks, 0, _keep=locs) for r in range(self.nranks)]

    def signal_pads(self) -> List[SymmBuffer]:
        return self.alloc(self.k.mpit_signal_words(), dtype=torch.int32)
