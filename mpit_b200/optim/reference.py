"""Reference-faithful optimizers in the torch-``optim`` calling convention.

``optim.<name>(opfunc, w, config, state=None) -> (w, [fx])`` exactly as the Lua originals:
``opfunc(w)`` returns ``(loss, grad)``; ``config['pclient']`` injects the parameter client
(asyncsgd/optim-downpour.lua:14, asyncsgd/optim-eamsgd.lua:17).  These run on plain torch
tensors (CPU workers over the host message layer, and they are the oracles the fused device
kernels are tested against).  Arithmetic is fp32, one rounding per tensor op, same op order as
the Lua source cited at each function.

``pclient`` duck type: ``reset(param, grad)``, ``async_recv_param()``, ``async_send_grad()``,
``async_send_param()``, ``ping()``, ``wait()``.
"""
from __future__ import annotations

import math
import time
from typing import Callable, Tuple

import torch

Opfunc = Callable[[torch.Tensor], Tuple[float, torch.Tensor]]


def _f32(x: float) -> float:
    """Narrow a host double to fp32 the way a Lua number is narrowed by a FloatTensor method."""
    return float(torch.tensor(x, dtype=torch.float32))


def _axpy(y: torch.Tensor, a: float, x: torch.Tensor) -> None:
    """y += a*x with a separate rounding for the product (TH cadd without FMA contraction)."""
    y.add_(x * _f32(a))


# ---------------------------------------------------------------------------------------------
# msgd -- asyncsgd/optim-msgd.lua:6-42 (+ BiCNN/optim-msgd.lua:22-24,46-48 when `single_push`)
# ---------------------------------------------------------------------------------------------
def msgd(opfunc: Opfunc, w: torch.Tensor, config: dict, state: dict = None):
    state = config if state is None else state
    lr = config.get("lr", 0.0)
    lrd = config.get("lrd", 0.0)
    lrp = config.get("lrp", 0.0)
    mom = config.get("mom", 0.0)
    mmax = config.get("mommax", 1.0)
    mlrd = config.get("momdecay", 0.0)
    l2wd = config.get("l2wd", 0.0)
    pc = config.get("pclient")
    single_push = config.get("single_push", False)
    state.setdefault("pversion", 0)
    if single_push and lrd != 0:  # BiCNN/optim-msgd.lua:22-24
        lr = lr / (1 + state["pversion"] * lrd)
    if mom > 0:
        if mlrd > 0:
            mom = min(mmax, 1 - 0.5 / (1 + state["pversion"] / mlrd))
        if "vt" not in state:
            state["vt"] = torch.zeros_like(w)
        state["vt"].mul_(_f32(mom))
        w.add_(state["vt"])
    fx, dfdx = opfunc(w)
    if l2wd != 0:
        _axpy(dfdx, l2wd, w)
    clr = lr
    if lrd > 0 and lrp > 0:
        clr = lr / math.pow(1 + state["pversion"] * lrd, lrp)
    _axpy(w, -clr, dfdx)
    if mom > 0:
        _axpy(state["vt"], -clr, dfdx)
    state["pversion"] += 1
    if single_push and pc is not None:
        pc.async_send_param()
        pc.wait()
    return w, [fx]


# ---------------------------------------------------------------------------------------------
# downpour -- asyncsgd/optim-downpour.lua:6-60
# ---------------------------------------------------------------------------------------------
def downpour(opfunc: Opfunc, w: torch.Tensor, config: dict, state: dict = None):
    state = config if state is None else state
    lr = config.get("lr", 0.0)
    lrd = config.get("lrd", 0.0)
    l2wd = config.get("l2wd", 0.0)
    pc = config.get("pclient")
    su = config.get("su", 0)
    state.setdefault("pversion", 0)
    state.setdefault("dusync", 0.0)
    if lrd != 0:
        lr = lr / (1 + state["pversion"] * lrd)
    fx, dfdx = opfunc(w)
    if l2wd != 0:
        _axpy(dfdx, l2wd, w)
    if pc is not None and su > 1:
        dfdx.mul_(_f32(-lr))
        if "dfdx" not in config:
            config["dfdx"] = torch.zeros_like(dfdx)
            pc.reset(w, config["dfdx"])
        config["dfdx"].add_(dfdx)
        if state["pversion"] % su == 0:
            pc.async_send_grad()
            pc.async_recv_param()
            t0 = time.perf_counter()
            pc.wait()
            state["dusync"] += time.perf_counter() - t0
            config["dfdx"].zero_()
        else:
            w.add_(dfdx)
    elif pc is not None and su == 1:
        dfdx.mul_(_f32(-lr))
        pc.async_send_grad()
        pc.async_recv_param()
        t0 = time.perf_counter()
        pc.wait()
        state["dusync"] += time.perf_counter() - t0
    else:
        raise AssertionError("downpour needs a pclient and su >= 1 (optim-downpour.lua:55-57)")
    state["pversion"] += 1
    return w, [fx]


# ---------------------------------------------------------------------------------------------
# eamsgd / easgd -- asyncsgd/optim-eamsgd.lua:7-79  (mom == 0 is EASGD)
# ---------------------------------------------------------------------------------------------
def eamsgd(opfunc: Opfunc, w: torch.Tensor, config: dict, state: dict = None):
    state = config if state is None else state
    lr = config.get("lr", 0.0)
    lrd = config.get("lrd", 0.0)
    lrp = config.get("lrp", 0.0)
    mom = config.get("mom", 0.0)
    l2wd = config.get("l2wd", 0.0)
    pc = config.get("pclient")
    mva = config.get("mva", 0.0)
    su = config.get("su", 1)
    state.setdefault("pversion", 0)
    state.setdefault("dusync", 0.0)
    out = {}

    def localupdate():
        if lr != 0:
            if mom > 0:
                if "vt" not in state:
                    state["vt"] = torch.zeros_like(w)
                state["vt"].mul_(_f32(mom))
                w.add_(state["vt"])
            fx, dfdx = opfunc(w)
            out["fx"] = fx
            if l2wd != 0:
                _axpy(dfdx, l2wd, w)
            clr = lr
            if lrd != 0 and lrp > 0:
                clr = lr / math.pow(1 + state["pversion"] * lrd, lrp)
            _axpy(w, -clr, dfdx)
            if mom > 0:
                _axpy(state["vt"], -clr, dfdx)
            state["pversion"] += 1

    if pc is not None and su > 0 and mva > 0:
        if state["pversion"] % su == 0:
            if "suw" not in config:
                config["suw"] = torch.zeros_like(w)
                config["sug"] = torch.zeros_like(w)
                pc.reset(config["suw"], config["sug"])
            pc.async_recv_param()
            t0 = time.perf_counter()
            pc.wait()
            state["dusync"] += time.perf_counter() - t0
            sug, suw = config["sug"], config["suw"]
            sug.copy_(w)
            sug.sub_(suw)
            sug.mul_(_f32(mva))
            pc.async_send_grad()
            t0 = time.perf_counter()
            pc.ping()
            state["dusync"] += time.perf_counter() - t0
            localupdate()
            w.sub_(sug)
        else:
            localupdate()
    else:
        raise AssertionError("eamsgd needs a pclient, su > 0 and mva > 0 (optim-eamsgd.lua:75-77)")
    return w, [out.get("fx")]


easgd = eamsgd  # mom == 0 (asyncsgd/optim-eamsgd.lua:4)


# ---------------------------------------------------------------------------------------------
# The five adaptive rules as pure tensor functions (shared by server-side and single variants)
# ---------------------------------------------------------------------------------------------
def rule_state(rule: str, p: torch.Tensor) -> dict:
    """State vectors a rule needs (BiCNN/pserver.lua:50-83)."""
    z = lambda: torch.zeros_like(p)
    if rule == "rmsprop":
        return {"gradAccum": z(), "gradSqAccum": z(), "update": z(), "t": 0}
    if rule == "adam":
        return {"m": z(), "v": z(), "t": 0}
    if rule == "adamax":
        return {"m": z(), "u": z(), "t": 0}
    if rule == "adagrad":
        return {"paramVariance": z(), "t": 0}
    if rule == "adadelta":
        return {"paramVariance": z(), "accDelta": z(), "t": 0}
    if rule == "add":
        return {"t": 0}
    raise ValueError(rule)


def apply_rule(rule: str, p: torch.Tensor, g: torch.Tensor, st: dict, hp: dict) -> None:
    """One update of ``p`` with gradient ``g``; ``st['t']`` counts updates (1-based after the call).

    rmsprop  BiCNN/pserver.lua:123-136  | adam     :140-155 (``step_div`` = stepDivAdam quirk)
    adamax   :156-171                   | adagrad  :172-183 | adadelta :184-195 | add :196-198
    """
    st["t"] += 1
    t = st["t"]
    if rule == "add":
        p.add_(g)
    elif rule == "rmsprop":
        decay, lr, mom, eps = hp["decay"], hp["lr"], hp["momentum"], hp["epsilon"]
        st["gradAccum"].mul_(_f32(decay)).add_(g * _f32(1 - decay))
        st["gradSqAccum"].mul_(_f32(decay)).add_((g * g) * _f32(1 - decay))
        rms = (st["gradSqAccum"] - st["gradAccum"] * st["gradAccum"]).add_(_f32(eps)).sqrt_()
        st["update"].mul_(_f32(mom)).add_((g / rms) * _f32(-lr))
        p.add_(st["update"])
    elif rule == "adam":
        b1, b2, lr, eps = hp["beta1"], hp["beta2"], hp["lr"], hp["epsilon"]
        sd = hp.get("step_div", 0)
        k = (t // sd + 1) if sd else t
        st["m"].mul_(_f32(b1)).add_(g * _f32(1 - b1))
        st["v"].mul_(_f32(b2)).add_((g * g) * _f32(1 - b2))
        den = st["v"].sqrt().add_(_f32(eps))
        lr_t = lr * math.sqrt(1 - b2 ** k) / (1 - b1 ** k)
        p.add_((st["m"] / den) * _f32(-lr_t))
    elif rule == "adamax":
        b1, b2, lr, eps = hp["beta1"], hp["beta2"], hp["lr"], hp["epsilon"]
        st["m"].mul_(_f32(b1)).add_(g * _f32(1 - b1))
        st["u"] = torch.maximum(st["u"] * _f32(b2), g.abs().add_(_f32(eps)))
        lr_t = lr / (1 - b1 ** t)
        p.add_((st["m"] / st["u"]) * _f32(-lr_t))
    elif rule == "adagrad":
        lr, lrd, eps = hp["lr"], hp.get("lrd", 0.0), hp["epsilon"]
        clr = lr / (1 + (t - 1) * lrd)
        st["paramVariance"].add_(g * g)
        std = st["paramVariance"].sqrt().add_(_f32(eps))
        p.add_((g / std) * _f32(-clr))
    elif rule == "adadelta":
        rho, eps, lr = hp.get("rho", 0.9), hp.get("epsilon", 1e-6), hp["lr"]
        st["paramVariance"].mul_(_f32(rho)).add_((g * g) * _f32(1 - rho))
        std = (st["paramVariance"] + _f32(eps)).sqrt_()
        delta = (st["accDelta"] + _f32(eps)).sqrt_().div_(std).mul_(g)
        p.add_(delta * _f32(-lr))
        st["accDelta"].mul_(_f32(rho)).add_((delta * delta) * _f32(1 - rho))
    else:
        raise ValueError(rule)


# ---------------------------------------------------------------------------------------------
# "parallel" clients of the server-side rules -- BiCNN/optim-{adam,adamax,adagrad,adadelta,rmsprop}.lua
# The client only ships raw gradients; the rule runs where the shard lives (BiCNN/pserver.lua:115-205).
# ---------------------------------------------------------------------------------------------
def _parallel(rule: str):
    def opt(opfunc: Opfunc, w: torch.Tensor, config: dict, state: dict = None):
        state = config if state is None else state
        pc = config.get("pclient")
        su = config.get("su", 0)
        mode = config.get("mode", "global")
        state.setdefault("pversion", 0)
        state.setdefault("dusync", 0.0)
        fx, dfdx = opfunc(w)

        def sync():
            pc.async_send_grad()
            pc.async_recv_param()
            t0 = time.perf_counter()
            pc.wait()
            state["dusync"] += time.perf_counter() - t0

        local_rms = rule == "rmsprop" and mode == "local"
        if local_rms:  # BiCNN/optim-rmsprop.lua:48-66,74-90: client computes the delta, server adds
            if "local_state" not in state:
                state["local_state"] = rule_state("rmsprop", w)
            ls = state["local_state"]
            scratch = torch.zeros_like(w)
            apply_rule("rmsprop", scratch, dfdx, ls, config)  # ls['update'] now holds this step's delta
        if pc is not None and su > 1:
            if "accumulated" not in config:
                config["accumulated"] = torch.zeros_like(dfdx)
                pc.reset(w, config["accumulated"])
            if local_rms:
                config["accumulated"].add_(state["local_state"]["update"])
                if state["pversion"] % su == 0:
                    sync()
                    config["accumulated"].zero_()
                else:
                    w.add_(state["local_state"]["update"])
            elif mode == "global":
                config["accumulated"].add_(dfdx)
                if state["pversion"] % su == 0:
                    sync()
                    config["accumulated"].zero_()
        elif pc is not None and su == 1:
            if local_rms:
                if not config.get("resetupdate"):
                    config["resetupdate"] = True
                    pc.reset(w, state["local_state"]["update"])
                sync()
            elif mode == "global":
                sync()
        else:
            raise AssertionError(f"{rule} (parallel) needs a pclient and su >= 1")
        state["pversion"] += 1
        return w, [fx]

    opt.__name__ = rule
    opt.__doc__ = f"Parallel client of the server-side {rule} rule (BiCNN/optim-{rule}.lua)."
    return opt


rmsprop = _parallel("rmsprop")
adam = _parallel("adam")
adamax = _parallel("adamax")
adagrad = _parallel("adagrad")
adadelta = _parallel("adadelta")


# ---------------------------------------------------------------------------------------------
# single-worker variants -- BiCNN/optim-*-single.lua: full rule on the worker, then push params
# ---------------------------------------------------------------------------------------------
def _single(rule: str):
    def opt(opfunc: Opfunc, w: torch.Tensor, config: dict, state: dict = None):
        state = config if state is None else state
        pc = config.get("pclient")
        state.setdefault("pversion", 0)
        fx, dfdx = opfunc(w)
        if "rule_state" not in state:
            state["rule_state"] = rule_state(rule, w)
        apply_rule(rule, w, dfdx, state["rule_state"], config)
        state["pversion"] += 1
        if pc is not None:
            pc.async_send_param()
            pc.wait()
        return w, [fx]

    opt.__name__ = rule + "single"
    opt.__doc__ = f"Single-worker {rule} (BiCNN/optim-{rule}-single.lua), pushes parameters after each step."
    return opt


rmspropsingle = _single("rmsprop")
adamsingle = _single("adam")
adamaxsingle = _single("adamax")
adagradsingle = _single("adagrad")
adadeltasingle = _single("adadelta")


def msgdsingle(opfunc, w, config, state=None):
    """BiCNN/optim-msgd.lua: Nesterov step, then push the whole parameter vector."""
    cfg = config
    cfg["single_push"] = True
    return msgd(opfunc, w, cfg, state)


OPTIMIZERS = {
    "sgd": msgd, "msgd": msgd, "downpour": downpour, "eamsgd": eamsgd, "easgd": easgd,
    "rmsprop": rmsprop, "adam": adam, "adamax": adamax, "adagrad": adagrad, "adadelta": adadelta,
    "rmspropsingle": rmspropsingle, "adamsingle": adamsingle, "adamaxsingle": adamaxsingle,
    "adagradsingle": adagradsingle, "adadeltasingle": adadeltasingle, "msgdsingle": msgdsingle,
}
