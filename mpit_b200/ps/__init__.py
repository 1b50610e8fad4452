"""Parameter server / client runtime, host path (reference: /root/reference/asyncsgd/init.lua,
pserver.lua, pclient.lua and the BiCNN copies with server-side optimizer rules).

The wire protocol is the reference's: eight message tags, zero-byte header/ack messages, one
coroutine per (client, server) pair, contiguous partition of the flat parameter vector with the
remainder on the last server, first-client seeding, and stop votes from every client.  It runs over
``mpit_b200.msg`` + ``mpit_b200.aio`` on CPU tensors: this is the plumbing path (CPU workers, config 1
of BASELINE.json) and the semantic oracle of the device path (``mpit_b200.parallel.ShardedCenter``),
where the same exchanges are single fused kernels over NVLink.

Deliberate differences from the reference:
* offsets are 0-based on the wire (the reference ships Lua's 1-based offsets, pclient.lua:118);
* the adadelta server rule is guarded by its own mode flag (the reference checks ``modeAdagrad``,
  BiCNN/pserver.lua:185);
* single-mode roles come from the configuration (``single_writer`` / ``single_readers``) instead of
  the hard-coded client ranks 0 and 2 (BiCNN/pserver.lua:280-289), and such servers do stop.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from .. import msg as mpiT
from ..aio import Queue, IOState, co_execute, co_ping, co_wait, aio_send, aio_recv, signal_INIT, signal_DONE
from ..optim.reference import apply_rule, rule_state
from ..parallel.center import partition

# message tags (/root/reference/asyncsgd/init.lua:3-10)
tag_ps_recv_init = 1
tag_ps_recv_grad = 2
tag_ps_send_param = 3
tag_ps_recv_param = 4
tag_ps_recv_header = 5
tag_ps_recv_stop = 6
tag_ps_recv_param_tail = 7
tag_ps_recv_grad_tail = 8

SERVER_RULES = ("rmsprop", "adam", "adamax", "adagrad", "adadelta")


def server_rule_config(opt: Optional[dict]):
    """Translate the launcher's option table (BiCNN/plaunch.lua:7-70 names) into (rule, hyper) for the
    shard owner; ``('add', {})`` when the server just accumulates (asyncsgd/pserver.lua:91)."""
    if not opt:
        return "add", {}
    name = opt.get("optimization", "")
    if name == "rmsprop" and opt.get("modeRMSProp", "global") == "global":
        return "rmsprop", dict(decay=opt["decayRMSProp"], lr=opt["lrRMSProp"], momentum=opt["momentumRMSProp"],
                               epsilon=opt["epsilonRMSProp"])
    if name == "adam" and opt.get("modeAdam", "global") == "global":
        return "adam", dict(lr=opt["lrAdam"], beta1=opt["beta1Adam"], beta2=opt["beta2Adam"],
                            epsilon=opt["epsilonAdam"], step_div=opt.get("stepDivAdam", 0))
    if name == "adamax" and opt.get("modeAdam", "global") == "global":
        return "adamax", dict(lr=opt["lrAdam"], beta1=opt["beta1Adam"], beta2=opt["beta2Adam"],
                              epsilon=opt["epsilonAdam"])
    if name == "adagrad" and opt.get("modeAdagrad", "global") == "global":
        return "adagrad", dict(lr=opt["lrAdagrad"], lrd=opt.get("lrDecayAdagrad", 0.0), epsilon=opt["epsilonAdagrad"])
    if name == "adadelta" and opt.get("modeAdadelta", "global") == "global":
        return "adadelta", dict(rho=opt.get("rhoAdadelta", 0.9), epsilon=opt.get("epsilonAdadelta", 1e-6),
                                lr=opt.get("lrAdadelta", 1.0))
    return "add", {}


class pServer:
    """One shard of the parameters and the service coroutines around it
    (/root/reference/asyncsgd/pserver.lua:12-168, BiCNN/pserver.lua)."""

    def __init__(self, conf: dict, state: Optional[IOState] = None):
        self.state = state if state is not None else IOState(io=False, on=False)
        self.rank = conf.get("rank", -1)
        self.cranks: List[int] = list(conf.get("cranks", []))
        self.mtype = mpiT.FLOAT
        self.mworld = conf.get("world", mpiT.COMM_WORLD)
        self.offset = -1
        self.size = -1
        self.p: Optional[torch.Tensor] = None
        self.g: Dict[int, torch.Tensor] = {}
        self.state["on"] = False
        self.state["io"] = False
        self.state["iostop"] = 0
        self.coq = Queue()
        self.conf = conf
        self.rule, self.hyper = server_rule_config(conf.get("opt"))
        self.rule_state = None
        self.n_updates = 0
        opt = conf.get("opt") or {}
        self.singlemode = bool(opt.get("singlemode", False))
        self.single_writer = opt.get("single_writer", self.cranks[0] if self.cranks else -1)
        self.single_readers = set(opt.get("single_readers", self.cranks[1:]))

    # -- coroutine bodies ------------------------------------------------------------------------
    def _recvinit(self, crank):
        yield signal_INIT
        cinfo = torch.zeros(2, dtype=torch.int64)
        yield from aio_recv(cinfo, 2, mpiT.LONG, crank, tag_ps_recv_init, self.mworld, self.state)
        off, size = int(cinfo[0]), int(cinfo[1])
        if self.offset == -1:
            self.offset, self.size = off, size
            # conf["device"] = "cuda:N": the shard lives on a GPU and the unmodified protocol runs GPU to GPU through
            # the device transport of the message layer (msg/device.py), like a CudaStorage handed to CUDA-aware MPI
            self.p = torch.zeros(size, dtype=torch.float32, device=self.conf.get("device", "cpu"))
            self.rule_state = rule_state(self.rule, self.p)   # BiCNN/pserver.lua:50-83
        else:
            assert self.offset == off and self.size == size, "clients disagree on this server's shard"
        self.g[crank] = torch.zeros(size, dtype=torch.float32, device=self.conf.get("device", "cpu"))   # one slot per client
        yield signal_DONE

    def _sendparam(self, crank):
        yield signal_INIT
        while self.state["on"]:
            yield from aio_recv(None, 0, self.mtype, crank, tag_ps_recv_header, self.mworld, self.state)
            if self.state["io"]:
                yield from aio_send(self.p, self.size, self.mtype, crank, tag_ps_send_param, self.mworld,
                                    self.state)
        yield signal_DONE

    def _recvgrad(self, crank):
        # no lock on p while gradients arrive: readers may see a partially updated shard, as in the
        # reference ("expect inconsistent read", asyncsgd/pserver.lua:81)
        yield signal_INIT
        while self.state["on"]:
            got = yield from aio_recv(self.g[crank], self.size, self.mtype, crank, tag_ps_recv_grad,
                                      self.mworld, self.state)
            if got:
                apply_rule(self.rule, self.p, self.g[crank], self.rule_state, self.hyper)
                self.n_updates += 1
            if self.state["on"]:
                yield from aio_send(None, 0, self.mtype, crank, tag_ps_recv_grad_tail, self.mworld, self.state)
        yield signal_DONE

    def _recvparam(self, crank, always=False):
        yield signal_INIT
        while True:
            yield from aio_recv(self.p, self.size, self.mtype, crank, tag_ps_recv_param, self.mworld, self.state)
            if self.state["on"]:
                yield from aio_send(None, 0, self.mtype, crank, tag_ps_recv_param_tail, self.mworld, self.state)
            if not always or not self.state["on"]:
                break
        yield signal_DONE

    def _recvstop(self, crank):
        yield signal_INIT
        tostop = torch.zeros(1, dtype=torch.uint8)
        got = yield from aio_recv(tostop, 1, mpiT.BYTE, crank, tag_ps_recv_stop, self.mworld, self.state)
        if got:
            self.state["iostop"] += 1
            if self.state["iostop"] == len(self.cranks):   # every client has voted
                self.state["on"] = False
                self.state["io"] = False
        yield signal_DONE

    # -- main loop -----------------------------------------------------------------------------------
    def start(self, usec: int = 0):
        """Phase A: learn the shard from every client; phase B: seed from the first client; then serve
        until every client has sent its stop vote (/root/reference/asyncsgd/pserver.lua:142-168)."""
        self.state["on"] = True
        self.state["io"] = True
        self.coq.clear()
        for crank in self.cranks:
            self.coq.push(co_execute(self._recvinit, (crank,)))
        co_wait(self.coq, usec)
        for i, crank in enumerate(self.cranks):
            if i == 0:
                self.coq.push(co_execute(self._recvparam, (crank,)))
                co_wait(self.coq, usec)
            self.coq.push(co_execute(self._recvstop, (crank,)))
            if self.singlemode:
                if crank == self.single_writer:
                    self.coq.push(co_execute(self._recvparam, (crank, True)))
                if crank in self.single_readers:
                    self.coq.push(co_execute(self._sendparam, (crank,)))
            else:
                self.coq.push(co_execute(self._recvgrad, (crank,)))
                self.coq.push(co_execute(self._sendparam, (crank,)))
        co_wait(self.coq, usec)


class pClient:
    """Client side (/root/reference/asyncsgd/pclient.lua:7-190)."""

    def __init__(self, conf: dict, state: Optional[IOState] = None):
        self.state = state if state is not None else IOState(io=False, on=False)
        self.rank = conf.get("rank", -1)
        self.sranks: List[int] = list(conf.get("sranks", []))
        self.cranks: List[int] = list(conf.get("cranks", []))
        self.pstorage: Optional[torch.Tensor] = conf.get("pstorage")
        self.gstorage: Optional[torch.Tensor] = conf.get("gstorage")
        self.plong = int(self.pstorage.numel()) if self.pstorage is not None else conf.get("plong", 0)
        self.sinfo: Dict[int, dict] = {}
        self.mtype = mpiT.FLOAT
        self.mworld = conf.get("world", mpiT.COMM_WORLD)
        self.coq = Queue()
        self.state["on"] = False
        self.state["io"] = False
        self.conf = conf

    # -- coroutine bodies --------------------------------------------------------------------------
    def _sendinit(self, srank, offset, size):
        yield signal_INIT
        self.sinfo[srank] = {"offset": offset, "size": size}
        cinfo = torch.tensor([offset, size], dtype=torch.int64)
        yield from aio_send(cinfo, 2, mpiT.LONG, srank, tag_ps_recv_init, self.mworld, self.state)
        yield signal_DONE

    def _sendstop(self, srank):
        yield signal_INIT
        tostop = torch.ones(1, dtype=torch.uint8)
        yield from aio_send(tostop, 1, mpiT.BYTE, srank, tag_ps_recv_stop, self.mworld, self.state)
        yield signal_DONE

    def _shard(self, t: torch.Tensor, srank) -> torch.Tensor:
        s = self.sinfo[srank]
        return t.view(-1)[s["offset"]:s["offset"] + s["size"]]   # alias, no copy (pclient.lua:51-53)

    def _sendgrad(self, grad, srank):
        yield signal_INIT
        sg = self._shard(grad, srank)
        yield from aio_send(sg, sg.numel(), self.mtype, srank, tag_ps_recv_grad, self.mworld, self.state)
        yield from aio_recv(None, 0, self.mtype, srank, tag_ps_recv_grad_tail, self.mworld, self.state)
        yield signal_DONE

    def _sendparam(self, param, srank):
        yield signal_INIT
        sp = self._shard(param, srank)
        yield from aio_send(sp, sp.numel(), self.mtype, srank, tag_ps_recv_param, self.mworld, self.state)
        yield from aio_recv(None, 0, self.mtype, srank, tag_ps_recv_param_tail, self.mworld, self.state)
        yield signal_DONE

    def _recvparam(self, param, srank):
        yield signal_INIT
        yield from aio_send(None, 0, self.mtype, srank, tag_ps_recv_header, self.mworld, self.state)
        sp = self._shard(param, srank)
        yield from aio_recv(sp, sp.numel(), self.mtype, srank, tag_ps_send_param, self.mworld, self.state)
        yield signal_DONE

    # -- fan-out: one coroutine per server (pclient.lua:87-114) ------------------------------------
    def async_recv_param(self):
        for srank in self.sranks:
            self.coq.push(co_execute(self._recvparam, (self.pstorage, srank)))

    def async_send_grad(self):
        for srank in self.sranks:
            self.coq.push(co_execute(self._sendgrad, (self.gstorage, srank)))

    def async_send_param(self, _ignored=None):
        for srank in self.sranks:
            self.coq.push(co_execute(self._sendparam, (self.pstorage, srank)))

    def _init(self):
        parts = partition(self.plong, len(self.sranks))   # pclient.lua:118-128
        for srank, (off, size) in zip(self.sranks, parts):
            self.coq.push(co_execute(self._sendinit, (srank, off, size)))
        co_wait(self.coq)
        if self.rank == self.cranks[0]:   # only the first client seeds the servers (pclient.lua:131-133)
            self.async_send_param()
        co_wait(self.coq)

    def ping(self, nb: Optional[int] = None):
        """Kick every queued coroutine once: the overlap hook (pclient.lua:139-144)."""
        nb = self.coq.len() if nb is None else nb
        for _ in range(nb):
            co_ping(self.coq)

    def reset(self, param: Optional[torch.Tensor], grad: Optional[torch.Tensor] = None):
        if param is not None:
            self.pstorage = param
            self.plong = int(param.numel())
            if grad is not None:
                self.gstorage = grad
                assert self.plong == int(grad.numel())

    def wait(self):
        co_wait(self.coq)

    def stop(self):
        self.wait()
        for srank in self.sranks:
            self.coq.push(co_execute(self._sendstop, (srank,)))
        self.wait()
        self.state["io"] = False
        self.state["on"] = False

    def start(self, param: Optional[torch.Tensor] = None, grad: Optional[torch.Tensor] = None):
        self.state["on"] = True
        self.state["io"] = True
        self.reset(param, grad)
        self._init()


def make_conf(rank: int, size: int, layout: str = "parity", world: int = mpiT.COMM_WORLD, **extra) -> dict:
    """Role tables of the launchers: ``parity`` = even ranks serve, odd ranks work (mlaunch.lua:40-46);
    ``halves`` = first half serve, second half work (ptest.lua:20-26)."""
    if layout == "parity":
        sranks = [i for i in range(size) if i % 2 == 0]
        cranks = [i for i in range(size) if i % 2 == 1]
    elif layout == "halves":
        sranks = list(range(size // 2))
        cranks = list(range(size // 2, size))
    else:
        raise ValueError(layout)
    conf = {"rank": rank, "world": world, "sranks": sranks, "cranks": cranks, "size": size}
    conf.update(extra)
    return conf
