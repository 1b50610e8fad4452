"""Device collectives over the symmetric heap: all-reduce (one-shot / two-shot, NVLS multimem when the
heap has a multicast mapping), broadcast, barrier and the all-reduce fused with the elastic pull.  (Two-sided Send/Recv on device buffers is the
device transport of the message layer, ``mpit_b200/msg/device.py``.)  Kernels: csrc/kernels/comm.cu.

Reference call sites: MPI_Allreduce / MPI_Iallreduce (test/testreduceall.lua:28, test/testireduceall.lua:32,
asyncsgd/testreduceall.lua:76), MPI_Bcast (wrapped only, mpifuncs.c:145), MPI_Send/Recv (test.lua:32-38).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from .._native import kernels, check, ptr, cur_stream
from .symm import SymmBuffer


class DeviceRequest:
    """Handle of a non-blocking device collective (MPI_Request for Iallreduce): a CUDA event recorded on
    the side stream that runs the kernel; ``test()`` polls it, ``wait()`` makes the current stream (and
    optionally the host) wait for it."""

    def __init__(self, stream: torch.cuda.Stream):
        self.event = torch.cuda.Event()
        self.event.record(stream)
        self.stream = stream

    def test(self) -> bool:
        return self.event.query()

    def wait(self, host: bool = False) -> None:
        torch.cuda.current_stream().wait_event(self.event)
        if host:
            self.event.synchronize()


class DeviceCollectives:
    """Collectives of one rank over buffers allocated from one symmetric heap.

    ``inp``/``out`` are symmetric buffers of ``capacity`` floats; the caller writes its contribution
    into ``inp.local`` (or passes a tensor to copy in) and reads the result from ``out.local``."""

    ONE_SHOT_MAX = 256 * 1024   # floats (1 MiB): below this the latency path wins on 2 ranks
    ONE_SHOT_MAX_NVLS = 16 * 1024   # from 3 ranks up the multimem two-shot is as fast at 64 KiB and faster beyond
                                    # (8 GPUs: 256 KiB 39.6 vs 102 us, 1 MiB 43.8 vs 52.6 us; profiles/ev_r2/n8/sweep.jsonl)

    def __init__(self, heap, capacity: int, multicast: bool = True):
        self.heap = heap
        self.k = kernels()
        self.capacity = capacity
        self.inp: SymmBuffer = heap.alloc(capacity, multicast=multicast)
        self.out: SymmBuffer = heap.alloc(capacity, multicast=multicast)
        # a pad of its own: the epoch words of the all-reduce must not be shared with a collective that may run
        # concurrently on another stream (Iallreduce next to a barrier of the training loop)
        self.sig: SymmBuffer = heap.alloc(self.k.mpit_signal_words(), dtype=torch.int32)
        self.flags: SymmBuffer = heap.alloc(int(self.k.mpit_allreduce_flag_words(capacity)), dtype=torch.int32)
        self.nranks, self.rank = self.inp.nranks, self.inp.rank
        self._in_ptrs, self._out_ptrs, self._sig_ptrs = self.inp.ptr_array(), self.out.ptr_array(), self.sig.ptr_array()
        self._flag_ptrs = self.flags.ptr_array()
        self._side = torch.cuda.Stream()

    def allreduce(self, x: Optional[torch.Tensor] = None, n: Optional[int] = None, *, average: bool = False,
                  algo: Optional[str] = None, elastic_x: Optional[torch.Tensor] = None, alpha: float = 0.0,
                  max_blocks: int = 0) -> torch.Tensor:
        """Sum (or mean) over ranks.  ``elastic_x`` fuses ``x -= alpha (x - mean)`` into the same kernel
        (synchronous EASGD: the all-reduce/average path of testreduceall fused with the elastic update).
        Returns a view of the result."""
        if x is not None:
            n = x.numel()
            self.inp.local[:n].copy_(x.reshape(-1))
        assert n is not None and n <= self.capacity
        if algo is None:
            algo = "oneshot" if n <= self._one_shot_max() else "twoshot"
        scale = 1.0 / self.nranks if average else 1.0
        two = algo == "twoshot"
        local_out = None if two else self.out.local
        # NVLS (multimem) pays from three ranks up and for the bandwidth algorithm only: measured on B200, 256 MiB
        # two-shot 624 us with multimem vs 771 us with peer loads on 8 GPUs but 700 vs 445 us on 2, and the 1 MiB
        # one-shot 105 us with multimem vs 50 us with peer loads on 8 GPUs (profiles/ar_tune_n8_r2_first.jsonl)
        use_mc = self._use_mc(two)
        check(self.k.mpit_allreduce(self._in_ptrs, self.inp.mc_ptr if use_mc else 0, self._out_ptrs if two else None,
                                    self.out.mc_ptr if (two and use_mc) else 0, self._sig_ptrs, self._flag_ptrs, self.nranks,
                                    self.rank, ptr(local_out), ptr(elastic_x), n, scale, alpha, 1 if two else 0,
                                    max_blocks, cur_stream()), "mpit_allreduce")
        return self.out.local[:n]

    def _one_shot_max(self) -> int:
        return self.ONE_SHOT_MAX_NVLS if (self.nranks > 2 and self.inp.mc_ptr != 0) else self.ONE_SHOT_MAX

    def _use_mc(self, two: bool) -> bool:
        use_mc = two and self.nranks > 2
        if os.environ.get("MPIT_AR_MC") in ("0", "1"):
            use_mc = os.environ["MPIT_AR_MC"] == "1"
        return bool(use_mc and self.inp.mc_ptr != 0 and self.out.mc_ptr != 0)

    def sync_easgd_step(self, x: torch.Tensor, v: Optional[torch.Tensor], wpk, wmap, *, alpha: float, mom: float = 0.0,
                        algo: Optional[str] = None, max_blocks: int = 0) -> None:
        """Boundary of a synchronous-EASGD step in one kernel: ``inp.local[:n]`` holds this rank's parameters after
        its local update (every rank wrote its own); computes the mean over ranks, ``x -= alpha (x - mean)``, the
        Nesterov look-ahead ``v *= mom; x += v`` (when ``v`` is given) and refreshes the operand copy ``wpk``."""
        n = x.numel()
        assert n <= self.capacity
        if algo is None:
            algo = "oneshot" if n <= self._one_shot_max() else "twoshot"
        two = algo == "twoshot"
        mc = self._use_mc(two)
        nmaps = 0 if wmap is None else max(1, int(wmap.numel()) // n)
        check(self.k.mpit_sync_easgd_step(self._in_ptrs, self.inp.mc_ptr if mc else 0, self._out_ptrs if two else None,
                                          self.out.mc_ptr if (two and mc) else 0, self._sig_ptrs, self._flag_ptrs,
                                          self.nranks, self.rank, ptr(x), ptr(v), ptr(wpk), ptr(wmap), nmaps, n, alpha,
                                          mom, 1 if two else 0, max_blocks, cur_stream()), "mpit_sync_easgd_step")

    def iallreduce(self, x: torch.Tensor, **kw) -> DeviceRequest:
        """MPI_Iallreduce: the same kernel on a side stream; Request = CUDA event."""
        self._side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self._side):
            self.allreduce(x, **kw)
        return DeviceRequest(self._side)

    def bcast(self, n: int, root: int, max_blocks: int = 0) -> torch.Tensor:
        """Broadcast ``out.local[:n]`` of ``root`` into every rank's ``out.local``."""
        check(self.k.mpit_bcast(self._out_ptrs, self.out.mc_ptr, self._sig_ptrs, self.nranks, self.rank, root, n,
                                max_blocks, cur_stream()), "mpit_bcast")
        return self.out.local[:n]

    def check_timeout(self) -> None:
        """Raise if a bounded device-side wait of one of this object's collectives gave up."""
        from .._native import NativeError
        err = int(self.sig.local[self.k.mpit_signal_words() - 64].item())
        if err != 0:
            raise NativeError("device collective timed out: a peer rank did not launch its kernel")

    def barrier(self) -> None:
        check(self.k.mpit_device_barrier(self._sig_ptrs, self.nranks, self.rank, cur_stream()), "mpit_device_barrier")
