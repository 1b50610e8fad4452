"""Device-side optimizers: fused kernel schedules around an engine's forward/backward.

An *engine* (e.g. ``models.cifar_cnn.CifarEngine``) exposes ``wpk`` (packed operand copy of the
parameters), ``gpk`` (packed gradients), ``wmap`` (flat -> packed index map) and ``fwd_bwd()``.
Each optimizer below owns the flat fp32 state in the reference's parameter order and drives one
training step as ``boundary kernel -> fwd_bwd``; the boundary kernel finishes step t-1 (apply),
does step t's exchange with the sharded centre over NVLink and prepares step t (Nesterov
look-ahead + operand refresh) in ONE pass over the parameters.  Steps are captured into CUDA
graphs after their first eager run.

Semantics mirror asyncsgd/optim-msgd.lua, optim-downpour.lua, optim-eamsgd.lua and the BiCNN
``*-single`` rules; see ``optim/reference.py`` for the line-by-line restatements they are tested
against.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch

from .. import ops
from .._native import launches
from ..parallel.center import ShardedCenter


class _GraphCache:
    """Eager on first sight of a step variant, capture on the second, replay afterwards."""

    def __init__(self, enabled: bool):
        self.enabled = enabled
        self.seen: Dict[tuple, int] = {}
        self.graphs: Dict[tuple, tuple] = {}

    def run(self, key: tuple, fn) -> None:
        if not self.enabled:
            fn()
            return
        if key in self.graphs:
            g, n = self.graphs[key]
            g.replay()
            launches.count += n
            return
        n_seen = self.seen.get(key, 0)
        self.seen[key] = n_seen + 1
        if n_seen == 0:
            fn()
            return
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        before = launches.count
        # thread_local: a native server thread (parallel/ps_device.py) may be launching and synchronising on its own
        # stream while this thread captures; only this thread's calls are checked against the capture
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            fn()
        n = launches.count - before
        launches.count = before
        self.graphs[key] = (g, n)
        g.replay()
        launches.count += n


class _DeviceOptBase:
    def __init__(self, engine, theta0: torch.Tensor, use_graphs: bool = True):
        self.engine = engine
        self.dev = engine.wpk.device
        self.x = theta0.detach().to(self.dev, torch.float32).clone().contiguous()
        self.n = self.x.numel()
        self.t = 0                      # pversion: number of completed local steps
        self.graphs = _GraphCache(use_graphs)
        self.pending_apply = False      # a gradient is sitting in engine.gpk, not yet applied

    def params(self) -> torch.Tensor:
        """Flat parameters with every computed gradient applied."""
        self.finalize()
        return self.x

    def finalize(self) -> None:  # overridden
        pass


class DeviceMSGD(_DeviceOptBase):
    """Nesterov momentum SGD on one GPU (asyncsgd/optim-msgd.lua:6-42)."""

    def __init__(self, engine, theta0, *, lr, mom=0.0, l2wd=0.0, lrd=0.0, lrp=0.0, mommax=1.0, momdecay=0.0,
                 use_graphs=True):
        super().__init__(engine, theta0, use_graphs and lrd == 0 and momdecay == 0)
        self.lr, self.mom, self.l2wd, self.lrd, self.lrp = lr, mom, l2wd, lrd, lrp
        self.mommax, self.momdecay = mommax, momdecay
        self.v = torch.zeros_like(self.x) if mom > 0 else None
        self.engine.load_flat(self.x)

    def _clr(self, t):
        if self.lrd > 0 and self.lrp > 0:
            return self.lr / math.pow(1 + t * self.lrd, self.lrp)
        return self.lr

    def _mom(self, t):
        if self.mom > 0 and self.momdecay > 0:
            return min(self.mommax, 1 - 0.5 / (1 + t / self.momdecay))
        return self.mom

    def _boundary(self, do_apply: bool, do_lookahead: bool):
        e = self.engine
        ops.local_step(self.x, self.v, e.gpk, e.gmap, None, e.wpk, e.wmap, clr=self._clr(self.t - 1),
                       l2wd=self.l2wd, mom_next=self._mom(self.t), do_apply=do_apply, use_mom=self.mom > 0,
                       do_lookahead=do_lookahead)

    def step(self) -> None:
        do_apply = self.pending_apply

        def body():
            self._boundary(do_apply, True)
            self.engine.fwd_bwd()

        self.graphs.run(("step", do_apply), body)
        self.pending_apply = True
        self.t += 1

    def finalize(self) -> None:
        if self.pending_apply:
            self._boundary(True, False)
            self.pending_apply = False


class DeviceEAMSGD(_DeviceOptBase):
    """Asynchronous EASGD/EAMSGD against an NVLink-sharded centre (asyncsgd/optim-eamsgd.lua:7-79).

    On a sync step ONE fused boundary kernel touches the parameters: it applies the previous gradient
    (``x -= clr g; v -= clr g; x -= d``), reads the centre shards from their owners, forms
    ``d = mva (x - centre)`` and accumulates it into the owners with ``red.add`` (fire and forget: it completes under
    the following forward/backward), then does the Nesterov look-ahead and refreshes the tensor-core operand copy.

    On the other steps nothing has to be exchanged, and the step is finished by the kernel that produces its last
    gradients anyway: the split-K reduction of the weight gradients applies each gradient element to its parameter,
    scales the momentum and writes the next look-ahead point into the operand copies (``fuse_reduce``; engines that
    offer ``opt_epi``; opt-in with ``fuse_reduce=True`` / ``MPIT_FUSE_REDUCE=1`` because it measured 3.6 % SLOWER on
    B200, see __init__).  With tau = 10 the boundary kernel then runs once in ten steps.  State convention between two
    such steps: ``x`` = parameters proper, ``v`` already scaled, operand copy = ``x + v`` (``_la_pending``); the float
    operations per parameter and their order are unchanged, so trajectories are bit-identical either way
    (tests/test_gpu_train_step.py)."""

    def __init__(self, engine, theta0, center: ShardedCenter, *, lr, mva, su=1, mom=0.0, l2wd=0.0, lrd=0.0,
                 lrp=0.0, seed_center: Optional[bool] = None, use_graphs=True, fuse_reduce: Optional[bool] = None):
        super().__init__(engine, theta0, use_graphs and not (lrd != 0 and lrp > 0))
        import os
        assert su > 0 and mva > 0, "eamsgd needs su > 0 and mva > 0 (optim-eamsgd.lua:50)"
        self.center = center
        self.lr, self.mva, self.su, self.mom, self.l2wd, self.lrd, self.lrp = lr, mva, su, mom, l2wd, lrd, lrp
        self.v = torch.zeros_like(self.x) if mom > 0 else None
        self.d = torch.zeros_like(self.x)
        self.prev_sync = False
        if fuse_reduce is None:
            # opt-in: measured on B200 it does NOT pay at this model size -- 0.1353 ms/step fused against 0.1305 with
            # the boundary kernel (tau = 10, batch 128): the boundary kernel hides under programmatic dependent launch
            # while the reduction, walking the packed order, pays for scattered x / v accesses on one thread in eight
            fuse_reduce = os.environ.get("MPIT_FUSE_REDUCE", "0") == "1"
        plan = getattr(engine, "_umma", None)
        self.fuse_reduce = bool(fuse_reduce and hasattr(engine, "opt_epi") and plan is not None
                                and getattr(plan, "wgrad", None) is not None
                                and getattr(plan.wgrad, "_red_opt", None) is not None and not plan.split_reduce)
        self._la_pending = False        # x = parameters proper, v scaled, operands at x + v (left by a fused reduction)
        self._operands_stale = False    # somebody re-packed the operands from x since (params() / test())
        # the first client seeds the servers (pclient.lua:131-133)
        if seed_center if seed_center is not None else center.rank == 0:
            center.put(self.x)
        self.engine.load_flat(self.x)

    def _clr(self, t):
        if self.lrd != 0 and self.lrp > 0:
            return self.lr / math.pow(1 + t * self.lrd, self.lrp)
        return self.lr

    def _boundary(self, do_apply, do_lookahead, do_sync, has_d_in, la_in=False):
        e = self.engine
        self.center.eamsgd_step(self.x, self.v, e.gpk, e.gmap, self.d, e.wpk, e.wmap, alpha=self.mva,
                                clr=self._clr(self.t - 1), l2wd=self.l2wd, mom_next=self.mom,
                                do_apply=do_apply, use_mom=self.mom > 0, do_lookahead=do_lookahead,
                                do_sync=do_sync, has_d_in=has_d_in, la_in=la_in)

    def step(self) -> None:
        do_apply = self.pending_apply
        do_sync = self.t % self.su == 0
        has_d_in = do_apply and self.prev_sync
        la_in = self._la_pending
        # the boundary kernel runs when there is something for it to do: a gradient to apply, an exchange, a first
        # look-ahead, or operands that were re-packed from x behind our back
        need_boundary = do_apply or do_sync or not la_in or self._operands_stale
        fuse = self.fuse_reduce and (self.t + 1) % self.su != 0
        epi = None
        if fuse:
            epi = dict(x=self.x, v=self.v, d=self.d, clr=self._clr(self.t), l2wd=self.l2wd, mom_next=self.mom,
                       use_mom=self.mom > 0, has_d=do_sync, la_in=la_in and not need_boundary)

        def body():
            if need_boundary:
                self._boundary(do_apply, True, do_sync, has_d_in, la_in)
            self.engine.opt_epi = epi
            try:
                self.engine.fwd_bwd()
            finally:
                self.engine.opt_epi = None

        self.graphs.run(("step", need_boundary, do_apply, do_sync, has_d_in, la_in, fuse), body)
        self._operands_stale = False
        if fuse:
            self.pending_apply, self.prev_sync, self._la_pending = False, False, True
        else:
            # without a boundary the state the step started from is still the state: x proper, v scaled
            self.pending_apply, self.prev_sync, self._la_pending = True, do_sync, la_in and not need_boundary
        self.t += 1

    def finalize(self) -> None:
        if self.pending_apply:
            self._boundary(True, False, False, self.prev_sync, self._la_pending)
            self.pending_apply = False
            self.prev_sync = False
            self._la_pending = False
        if self._la_pending:
            self._operands_stale = True     # callers of params() may re-pack the operands from x (test, load_flat)


class DeviceSyncEASGD(_DeviceOptBase):
    """Synchronous EASGD / EAMSGD: every ``su`` steps all workers meet, the centre is the mean of their parameters and
    each worker moves toward it, ``x_i -= alpha (x_i - mean)`` -- the all-reduce/average path of
    asyncsgd/testreduceall.lua:72-77 (MPI_Allreduce, test/testreduceall.lua:28) used as a training path and fused with
    the elastic update of optim-eamsgd.lua:62-70.

    A sync step is two kernels: the local update of step t-1 (which also drops the parameters into the symmetric
    input buffer), then ONE kernel that all-reduces over NVLink (NVLS multimem from three ranks up), applies the
    elastic update, the Nesterov look-ahead of step t and the operand refresh.  Other steps are the single fused
    boundary kernel of :class:`DeviceMSGD`.  All ranks must step in lock-step."""

    def __init__(self, engine, theta0, heap, *, lr, mva, su=1, mom=0.0, l2wd=0.0, use_graphs=True):
        super().__init__(engine, theta0, use_graphs)
        from ..parallel.collectives import DeviceCollectives
        assert su > 0 and mva > 0
        self.lr, self.mva, self.su, self.mom, self.l2wd = lr, mva, su, mom, l2wd
        self.coll = DeviceCollectives(heap, self.n)
        self.v = torch.zeros_like(self.x) if mom > 0 else None
        self.engine.load_flat(self.x)

    def _boundary(self, do_apply: bool, do_lookahead: bool, do_sync: bool) -> None:
        e = self.engine
        if do_sync:
            # apply(t-1); the same pass writes x into the symmetric input buffer (identity "operand copy")
            ops.local_step(self.x, self.v, e.gpk, e.gmap, None, self.coll.inp.local, None, clr=self.lr, l2wd=self.l2wd,
                           mom_next=0.0, do_apply=do_apply, use_mom=self.mom > 0, do_lookahead=False)
            self.coll.sync_easgd_step(self.x, self.v if do_lookahead else None, e.wpk, e.wmap, alpha=self.mva,
                                      mom=self.mom)
        else:
            ops.local_step(self.x, self.v, e.gpk, e.gmap, None, e.wpk, e.wmap, clr=self.lr, l2wd=self.l2wd,
                           mom_next=self.mom, do_apply=do_apply, use_mom=self.mom > 0, do_lookahead=do_lookahead)

    def step(self) -> None:
        do_apply = self.pending_apply
        do_sync = self.t % self.su == 0

        def body():
            self._boundary(do_apply, True, do_sync)
            self.engine.fwd_bwd()

        self.graphs.run(("step", do_apply, do_sync), body)
        self.pending_apply = True
        self.t += 1

    def finalize(self) -> None:
        if self.pending_apply:
            self._boundary(True, False, False)
            self.pending_apply = False


class DeviceDownpour(_DeviceOptBase):
    """Downpour / hogwild SGD against the sharded centre (asyncsgd/optim-downpour.lua:6-60).

    After the backward pass ONE kernel finishes the step: ``g' = -lr (g + l2wd x)``, ``acc += g'``; on a sync step
    (every ``su``-th, step 0 included) the accumulator goes into the owners' shards with a returning accumulate over
    NVLink and the worker continues from the value found there plus its own contribution (push followed by pull, one
    round trip), the accumulator is cleared; on the other steps the worker moves on its own.  The tensor-core operand
    copies are refreshed in the same pass."""

    def __init__(self, engine, theta0, center: ShardedCenter, *, lr, su=1, l2wd=0.0, lrd=0.0,
                 seed_center: Optional[bool] = None, use_graphs=True):
        super().__init__(engine, theta0, use_graphs and lrd == 0)
        assert su >= 1
        self.center = center
        self.lr, self.su, self.l2wd, self.lrd = lr, su, l2wd, lrd
        self.acc = torch.zeros_like(self.x)
        if seed_center if seed_center is not None else center.rank == 0:
            center.put(self.x)
        self.engine.load_flat(self.x)

    def boundary(self, do_sync: bool, lr: float) -> None:
        from .._native import kernels, check, ptr, cur_stream
        e, c = self.engine, self.center
        wmap = getattr(e, "wmap", None)
        nmaps = 0 if wmap is None else max(1, int(wmap.numel()) // self.n)
        check(kernels().mpit_downpour_fused_step(c._ptrs, c.nranks, c.rank, ptr(self.x), ptr(self.acc), ptr(e.gpk),
                                                 ptr(getattr(e, "gmap", None)), ptr(e.wpk), ptr(wmap), nmaps, self.n,
                                                 -lr, self.l2wd, int(do_sync), cur_stream()),
              "mpit_downpour_fused_step")

    def step(self) -> None:
        lr = self.lr / (1 + self.t * self.lrd) if self.lrd != 0 else self.lr
        do_sync = self.t % self.su == 0

        def body():
            self.engine.fwd_bwd()
            self.boundary(do_sync, lr)

        self.graphs.run(("step", do_sync), body)
        self.t += 1


class DeviceRule(_DeviceOptBase):
    """Single-worker adaptive rules (BiCNN/optim-*-single.lua) fused into one pass, optionally
    pushing the parameters to the centre afterwards (``async_send_param``)."""

    def __init__(self, engine, theta0, rule: str, hyper: dict, center: Optional[ShardedCenter] = None, *,
                 l1=0.0, l2=0.0, clip=0.0, use_graphs=False):
        super().__init__(engine, theta0, False)
        self.rule, self.hyper, self.center = rule, dict(hyper), center
        self.l1, self.l2, self.clip = l1, l2, clip
        z = lambda: torch.zeros_like(self.x)
        self.s0 = z() if rule != "add" else None
        self.s1 = z() if rule in ("rmsprop", "adam", "adamax", "adadelta") else None
        self.s2 = z() if rule == "rmsprop" else None
        self.engine.load_flat(self.x)

    def step(self) -> None:
        e = self.engine
        e.fwd_bwd()
        self.t += 1
        h = ops.rule_hyper(self.rule, self.t, **self.hyper)
        ops.rule_step(self.rule, self.x, e.gpk, h, self.s0, self.s1, self.s2, gmap=e.gmap, wpk=e.wpk,
                      wmap=e.wmap, l1=self.l1, l2=self.l2, clip=self.clip)
        if self.center is not None:
            self.center.put(self.x)


class DeviceRuleParallel(_DeviceOptBase):
    """Server-side adaptive rules on the device path (BiCNN/optim-{adam,adamax,adagrad,adadelta,rmsprop}.lua
    clients + BiCNN/pserver.lua:115-205 server rules), asynchronous like the reference.

    Every GPU is a client *and* the owner of one parameter shard together with that shard's optimizer state (the
    moments live only with the owner, as on the reference's servers).  A sync step is three kernels and no barrier:
    ``push`` (gradient prologue, su-step accumulation, stores into this client's staging slot on every owner -- the
    per-client ``storage.g[crank]`` of pserver.lua:49-58 -- and an arrival ticket per owner), ``drain`` (the owner
    applies the gradients that have ARRIVED, one client at a time in arrival order, and never waits for the rest) and
    ``pull`` (fresh parameters + operand copies).  Update counts and step sizes live in device memory, so the whole
    step is one CUDA graph.  ``server="thread"`` moves the draining to a native host thread with its own stream
    (the reference's dedicated server process); ``"inline"`` drains once per step on the training stream.
    """

    def __init__(self, engine, theta0, rule: str, hyper: dict, center: ShardedCenter, heap, *, su: int = 1,
                 l1=0.0, l2=0.0, clip=0.0, gscale: float = 1.0, seed_center: Optional[bool] = None,
                 server: str = "inline", server_interval_us: int = 20, use_graphs: bool = True):
        super().__init__(engine, theta0, use_graphs)
        from ..parallel.ps_device import AsyncDevicePS
        assert server in ("inline", "thread")
        self.rule, self.hyper, self.center, self.su, self.server = rule, dict(hyper), center, su, server
        self.l1, self.l2, self.clip, self.gscale = l1, l2, clip, gscale    # gscale = -lr: Downpour's push of -lr*g
        self.nranks, self.rank = center.nranks, center.rank
        self.ps = AsyncDevicePS(heap, self.n, rule, hyper, center)
        self.acc = torch.zeros_like(self.x) if su > 1 else None     # config.accumulated (optim-adam.lua:27-30)
        if seed_center if seed_center is not None else center.rank == 0:
            center.put(self.x)
        self.engine.load_flat(self.x)
        if server == "thread":
            self.ps.start_server(server_interval_us)

    def _sources(self):
        e = self.engine
        gmap = getattr(e, "gmap", None)
        if gmap is None:
            return dict(gflat=e.gpk, gpk=None, gmap=None, map_base=0)
        return dict(gflat=getattr(e, "gflat", None), gpk=e.gpk, gmap=gmap, map_base=getattr(e, "map_base", 0))

    def exchange(self, do_sync: bool, active=None, reg_loss=None) -> None:
        """push [-> drain] -> pull on the current stream (what follows the backward pass of a step)."""
        e = self.engine
        self.ps.push(x=self.x, acc=self.acc, active=active, reg_loss=reg_loss, l1=self.l1, l2=self.l2, clip=self.clip,
                     gscale=self.gscale, do_push=do_sync, **self._sources())
        if do_sync:
            if self.server == "inline":
                self.ps.drain()
            self.ps.pull(self.x, e.wpk, getattr(e, "wmap", None), getattr(e, "map_base", 0))

    def step(self, fwd_bwd=None) -> None:
        do_sync = self.t % self.su == 0        # pversion % su == 0, step 0 included (optim-adam.lua:32)
        run = fwd_bwd if fwd_bwd is not None else self.engine.fwd_bwd

        def body():
            run()
            self.exchange(do_sync)

        self.graphs.run(("step", do_sync), body)
        self.t += 1

    def close(self) -> int:
        """Stops the server thread (if any); returns the number of drains it ran."""
        return self.ps.stop_server()
