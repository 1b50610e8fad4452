"""Asynchronous parameter server with server-side update rules on the NVLink device path.

Reference: the BiCNN servers (``/root/reference/BiCNN/pserver.lua:115-205``) keep one receive buffer per client
(``storage.g[crank]``, ``:49-58``), apply rmsprop / adam / adamax / adagrad / adadelta once per incoming gradient in
arrival order and never wait for a client that has not pushed; the clients
(``/root/reference/BiCNN/optim-adam.lua:24-57`` and siblings) push a raw gradient, pull the parameters and go on.

Here every GPU is a client *and* the owner of shard ``rank`` of the parameters together with that shard's rule
state.  The three operations are kernels of ``csrc/kernels/ps_async.cu``:

* :meth:`push` -- gradient prologue + stores into this client's staging slot on every owner + arrival tickets;
* :meth:`drain` -- the owner applies whatever has arrived, one client at a time in arrival order (update count and
  step sizes live in device memory, so the call is CUDA-graph capturable);
* :meth:`pull` -- one-sided read of all shards, refreshing the tensor-core operand copies in the same pass.

``start_server`` hands the draining to a native host thread with its own high-priority stream, the counterpart of the
reference's dedicated server processes: a rank that computes slowly then no longer delays the gradients that other
ranks send to its shard.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from .._native import kernels, check, ptr, cur_stream, NativeError
from ..ops.optim_ops import RULES, rule_hyper
from .center import ShardedCenter, shard_capacity


def _hyper10(rule: str, hyper: dict):
    """{a, b, c, d, eps} as the kernels want them plus the doubles the device needs for the per-update step size."""
    h6 = rule_hyper(rule, 1, **hyper) if rule != "add" else [0.0] * 6
    lr = float(hyper.get("lr", 0.0))
    vals = [h6[0], h6[1], h6[2], h6[3], h6[4], lr, float(hyper.get("beta1", 0.0)), float(hyper.get("beta2", 0.0)),
            float(hyper.get("lrd", 0.0)), float(hyper.get("step_div", 0))]
    return (ctypes.c_double * 10)(*vals)


class AsyncDevicePS:
    def __init__(self, heap, n: int, rule: str, hyper: dict, center: Optional[ShardedCenter] = None,
                 slots=None, ctl=None):
        self.k = kernels()
        self.heap, self.n, self.rule = heap, n, rule
        self.center = center if center is not None else ShardedCenter.create(heap, n)
        self.nranks, self.rank = self.center.nranks, self.center.rank
        self.cap = shard_capacity(n, self.nranks)
        self.dev = self.center.buf.local.device
        # [client][shard element] on every owner, and the control block
        self.slots = slots if slots is not None else heap.alloc(self.cap * self.nranks)
        self.ctl = ctl if ctl is not None else heap.alloc(int(self.k.mpit_ps_ctl_words()), dtype=torch.int32)
        self._slot_ptrs = self.slots.offset_ptr_array(4 * self.cap * self.rank)   # my slot inside each owner
        self._ctl_ptrs = self.ctl.ptr_array()
        self.shard_size = self.center.parts[self.rank][1]
        z = lambda need: torch.zeros(self.cap, device=self.dev) if need else None
        self.s0 = z(rule != "add")
        self.s1 = z(rule in ("rmsprop", "adam", "adamax", "adadelta"))
        self.s2 = z(rule == "rmsprop")
        self._rule_id = RULES[rule]
        self._hyper = _hyper10(rule, hyper)
        self._server = None

    @classmethod
    def local_group(cls, heap, n: int, rule: str, hyper: dict):
        """One instance per 'rank' of a single-process :class:`LocalPeerHeap` (tests, one-GPU runs)."""
        k = kernels()
        S = heap.nranks
        cap = shard_capacity(n, S)
        centers = [ShardedCenter(b, n) for b in heap.alloc(cap)]
        slots = heap.alloc(cap * S)
        ctls = heap.alloc(int(k.mpit_ps_ctl_words()), dtype=torch.int32)
        return [cls(heap, n, rule, hyper, centers[r], slots[r], ctls[r]) for r in range(S)]

    # -- client side -----------------------------------------------------------------------------------------
    def push(self, *, gflat=None, gpk=None, gmap=None, map_base: int = 0, x=None, acc=None, active=None,
             reg_loss=None, l1=0.0, l2=0.0, clip=0.0, gscale=1.0, do_push=True) -> None:
        check(self.k.mpit_ps_push(self._slot_ptrs, self._ctl_ptrs, self.nranks, self.rank, ptr(gflat), ptr(gpk),
                                  ptr(gmap), map_base, ptr(x), ptr(acc), ptr(active), ptr(reg_loss), self.n, l1, l2,
                                  clip, gscale, int(do_push), cur_stream()), "mpit_ps_push")

    def pull(self, dst: torch.Tensor, wpk=None, wmap=None, map_base: int = 0) -> None:
        self.center.pull(dst, wpk, wmap, map_base)

    # -- owner side ------------------------------------------------------------------------------------------
    def drain(self, max_take: int = 0) -> None:
        check(self.k.mpit_ps_drain(self._ctl_ptrs, self.nranks, self.rank, self._rule_id, self._hyper,
                                   ptr(self.center.buf.local), ptr(self.s0), ptr(self.s1), ptr(self.s2),
                                   ptr(self.slots.local), self.cap, self.shard_size, max_take, cur_stream()),
              "mpit_ps_drain")

    def start_server(self, interval_us: int = 20) -> None:
        """Drain from a native thread on a high-priority stream until :meth:`stop_server`."""
        if self._server is not None:
            return
        h = ctypes.c_void_p(0)
        check(self.k.mpit_ps_server_start(self.dev.index or 0, self._ctl_ptrs, self.nranks, self.rank, self._rule_id,
                                          self._hyper, ptr(self.center.buf.local), ptr(self.s0), ptr(self.s1),
                                          ptr(self.s2), ptr(self.slots.local), self.cap, self.shard_size, interval_us,
                                          ctypes.byref(h)), "mpit_ps_server_start")
        self._server = h

    def stop_server(self) -> int:
        """Returns the number of drain rounds the thread ran."""
        if self._server is None:
            return 0
        rounds = ctypes.c_longlong(0)
        rc = self.k.mpit_ps_server_stop(self._server, ctypes.byref(rounds))
        self._server = None
        if rc != 0:
            raise NativeError(f"parameter-server thread failed with code {rc}")
        return int(rounds.value)

    # -- monitoring --------------------------------------------------------------------------------------------
    def _counter(self, which: int) -> int:
        out = ctypes.c_uint(0)
        rc = self.k.mpit_ps_counter(ptr(self.ctl.local), which, ctypes.byref(out))
        if rc != 0:
            raise NativeError(f"mpit_ps_counter failed: {rc}")
        return int(out.value)

    @property
    def applied(self) -> int:
        """Updates applied to my shard so far (the server's adam_t / pversion)."""
        return self._counter(0)

    @property
    def pushes(self) -> int:
        return self._counter(1)

    @property
    def arrivals(self) -> int:
        return self._counter(3)

    @property
    def max_wait_us(self) -> float:
        """Longest time a push of mine waited for an owner to free its slot (flow control), at the nominal 1.9 GHz."""
        return self._counter(4) * 1024 / 1.9e3

    def check_timeout(self) -> None:
        e = self._counter(2)
        if e != 0:
            raise NativeError(f"asynchronous parameter server: a push waited more than 2 s for owner {e - 1} to "
                              f"acknowledge the previous one (that owner never drained)")

    def __del__(self):
        try:
            self.stop_server()
        except Exception:
            pass
