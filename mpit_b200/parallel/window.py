"""MPI one-sided communication on the symmetric heap: the native API of the NVLink transport.

The reference wraps ``MPI_Win_create / Put / Get / Accumulate / Win_fence / Win_lock / Win_unlock``
(/root/reference/mpifuncs.c:2337, 1656, 1131, 9, 2367, 2424, 2485) without ever calling them; its parameter
server re-implements the same three verbs with tagged two-sided messages (pull = Get, push-grad = Accumulate,
seed = Put: /root/reference/asyncsgd/pclient.lua:49-84).  Here the window *is* the hardware model: every rank's
window memory is mapped into every other rank, ``Put`` is ``st.sys``, ``Get`` is ``ld.sys``, ``Accumulate(SUM)`` is
``red.add.sys`` issued from a kernel of the origin GPU; no process on the target takes part (true passive target).

Synchronisation:
* ``fence()``          -- MPI_Win_fence: device-side barrier over all ranks (active target).
* ``lock()/unlock()``  -- MPI_Win_lock/unlock: passive target.  NVLink atomics need no lock for Accumulate;
                          ``unlock`` only waits until this rank's operations have left the GPU (stream sync), which
                          is MPI's completion guarantee at the origin.  An exclusive lock for read-modify-write
                          sequences is not provided (the reference's servers add without locks, pserver.lua:81).
"""
from __future__ import annotations

from typing import Optional

import torch

from .._native import check, cur_stream, kernels
from .symm import SymmBuffer

PUT, GET, ACC = 0, 1, 2
LOCK_EXCLUSIVE, LOCK_SHARED = 234, 235     # the MPI constants' usual values (lua-mpi.h:165-175 exports them)


class DeviceWindow:
    """``Win_create`` over ``numel`` floats of symmetric memory per rank."""

    def __init__(self, heap, numel: int, buf: Optional[SymmBuffer] = None, sig: Optional[SymmBuffer] = None):
        self.k = kernels()
        self.buf: SymmBuffer = buf if buf is not None else heap.alloc(numel)
        self.sig: SymmBuffer = sig if sig is not None else heap.signal_pad()
        self.numel = numel
        self.rank, self.nranks = self.buf.rank, self.buf.nranks
        self._sig_ptrs = self.sig.ptr_array()

    @property
    def local(self) -> torch.Tensor:
        """This rank's window memory."""
        return self.buf.local[:self.numel]

    def _rma(self, op: int, origin: torch.Tensor, target_rank: int, target_disp: int, count: Optional[int],
             scale: float) -> None:
        n = origin.numel() if count is None else count
        if not (0 <= target_rank < self.nranks):
            raise ValueError(f"target rank {target_rank} outside 0..{self.nranks - 1}")
        if target_disp < 0 or target_disp + n > self.numel or n > origin.numel():
            raise ValueError(f"window access [{target_disp}, {target_disp + n}) outside 0..{self.numel}")
        if origin.dtype != torch.float32 or not origin.is_cuda or not origin.is_contiguous():
            raise TypeError("origin must be a contiguous float32 CUDA tensor")
        remote = self.buf.ptrs[target_rank] + 4 * target_disp
        check(self.k.mpit_peer_rma(op, remote, origin.data_ptr(), n, scale, cur_stream()), "mpit_peer_rma")

    def put(self, origin: torch.Tensor, target_rank: int, target_disp: int = 0, count: Optional[int] = None) -> None:
        """MPI_Put: window[target][disp : disp+n] = origin."""
        self._rma(PUT, origin, target_rank, target_disp, count, 1.0)

    def get(self, origin: torch.Tensor, target_rank: int, target_disp: int = 0, count: Optional[int] = None) -> None:
        """MPI_Get: origin = window[target][disp : disp+n]."""
        self._rma(GET, origin, target_rank, target_disp, count, 1.0)

    def accumulate(self, origin: torch.Tensor, target_rank: int, target_disp: int = 0, count: Optional[int] = None,
                   scale: float = 1.0) -> None:
        """MPI_Accumulate with MPI_SUM: window[target][disp : disp+n] += scale * origin (atomic per element)."""
        self._rma(ACC, origin, target_rank, target_disp, count, scale)

    def fence(self) -> None:
        """MPI_Win_fence: every rank's preceding operations are complete and visible afterwards."""
        check(self.k.mpit_device_barrier(self._sig_ptrs, self.nranks, self.rank, cur_stream()), "mpit_device_barrier")

    def lock(self, target_rank: int, lock_type: int = LOCK_SHARED) -> None:
        """MPI_Win_lock (shared): nothing to acquire -- see the module docstring."""
        if lock_type == LOCK_EXCLUSIVE:
            raise NotImplementedError("exclusive window locks are not provided (Accumulate is atomic without one)")

    def unlock(self, target_rank: int) -> None:
        """MPI_Win_unlock: this rank's operations on the window have completed at the origin."""
        torch.cuda.current_stream().synchronize()

    flush = unlock

    def free(self) -> None:
        self.buf = None
