"""Deterministic single-worker trajectories of the reference-convention optimizers vs an independent NumPy
restatement of the Lua sources (asyncsgd/optim-msgd.lua, optim-downpour.lua, optim-eamsgd.lua,
BiCNN/optim-*-single.lua, BiCNN/pserver.lua rules)."""
import math

import numpy as np
import pytest
import torch

from mpit_b200.optim import reference as R

f32 = np.float32


class LocalPS:
    """pclient + plain-add pserver in one object (asyncsgd/pserver.lua:91), single worker."""

    def __init__(self, center):
        self.center = center.clone()
        self.sent_params = 0

    def reset(self, p, g):
        self.p, self.g = p, g

    def async_recv_param(self):
        self.p.copy_(self.center)

    def async_send_grad(self):
        self.center.add_(self.g)

    def async_send_param(self, _=None):
        self.center.copy_(self.p)
        self.sent_params += 1

    def ping(self):
        pass

    def wait(self):
        pass


def _grads(k, n=257, seed=0):
    rng = np.random.RandomState(seed)
    return [rng.randn(n).astype(f32) * f32(0.1) for _ in range(k)]


def test_msgd_nesterov_with_schedules():
    gs = _grads(8)
    w_np = np.linspace(-1, 1, 257).astype(f32)
    w = torch.from_numpy(w_np.copy())
    cfg = {"lr": 0.05, "lrd": 0.01, "lrp": 0.75, "mom": 0.9, "mommax": 0.95, "momdecay": 3.0, "l2wd": 1e-3}
    it = iter(gs)
    for _ in gs:
        R.msgd(lambda x: (0.0, torch.from_numpy(next(it).copy())), w, cfg)
    v = np.zeros_like(w_np)
    for t, g in enumerate(gs):          # optim-msgd.lua:20-39
        mom = min(0.95, 1 - 0.5 / (1 + t / 3.0))
        v = v * f32(mom)
        w_np = w_np + v
        g = g + w_np * f32(1e-3)
        clr = f32(-(0.05 / math.pow(1 + t * 0.01, 0.75)))
        w_np = w_np + g * clr
        v = v + g * clr
    assert np.array_equal(w.numpy(), w_np)
    assert cfg["pversion"] == 8


def test_plain_sgd_when_mom_missing_is_the_reference_quirk():
    """goot.lua:112-115 passes mommax instead of mom: optim.msgd then runs momentum-free SGD."""
    w = torch.ones(5)
    cfg = {"lr": 0.1, "mommax": 0.99}
    R.msgd(lambda x: (0.0, torch.ones(5)), w, cfg)
    assert "vt" not in cfg and torch.allclose(w, torch.full((5,), 0.9))


@pytest.mark.parametrize("su", [1, 3])
def test_downpour_single_worker(su):
    gs = _grads(7, seed=1)
    w0 = np.linspace(0, 1, 257).astype(f32)
    w = torch.from_numpy(w0.copy())
    ps = LocalPS(torch.from_numpy(w0.copy()))
    gbuf = torch.zeros(257)
    cfg = {"lr": 0.1, "su": su, "pclient": ps}
    if su == 1:
        ps.reset(w, gbuf)
    it = iter(gs)

    def op(x):
        g = torch.from_numpy(next(it).copy())
        if su == 1:
            gbuf.copy_(g)
            return 0.0, gbuf
        return 0.0, g

    for _ in gs:
        R.downpour(op, w, cfg)
    # restatement (optim-downpour.lua:26-54)
    x, c, acc = w0.copy(), w0.copy(), np.zeros(257, f32)
    for t, g in enumerate(gs):
        d = g * f32(-0.1)
        if su == 1:
            c = c + d
            x = c.copy()
        else:
            acc = acc + d
            if t % su == 0:          # step 0 is a sync step
                c = c + acc
                x = c.copy()         # the pull OVERWRITES the local parameters
                acc[:] = 0
            else:
                x = x + d
    assert np.array_equal(w.numpy(), x)
    assert np.array_equal(ps.center.numpy(), c)


@pytest.mark.parametrize("su,mom", [(1, 0.0), (2, 0.9), (4, 0.99)])
def test_eamsgd_single_worker(su, mom):
    gs = _grads(9, seed=2)
    w0 = np.linspace(-1, 1, 257).astype(f32)
    c0 = np.zeros(257, f32)
    w = torch.from_numpy(w0.copy())
    ps = LocalPS(torch.from_numpy(c0.copy()))
    cfg = {"lr": 0.05, "mom": mom, "mva": 0.25, "su": su, "pclient": ps}
    it = iter(gs)
    for _ in gs:
        R.eamsgd(lambda x: (0.0, torch.from_numpy(next(it).copy())), w, cfg)
    x, c, v = w0.copy(), c0.copy(), np.zeros(257, f32)
    for t, g in enumerate(gs):          # optim-eamsgd.lua:26-47,50-74
        sync = t % su == 0
        if sync:
            d = (x - c) * f32(0.25)
            c = c + d
        if mom > 0:
            v = v * f32(mom)
            x = x + v
        x = x + g * f32(-0.05)
        if mom > 0:
            v = v + g * f32(-0.05)
        if sync:
            x = x - d
    assert np.array_equal(w.numpy(), x)
    assert np.array_equal(ps.center.numpy(), c)


def test_eamsgd_requires_client_and_rates():
    with pytest.raises(AssertionError):
        R.eamsgd(lambda x: (0.0, x), torch.zeros(3), {"lr": 0.1, "mva": 0.1})     # optim-eamsgd.lua:75-77


def _np_rule(rule, p, g, st, hp, t):
    """BiCNN/pserver.lua:123-195 / optim-*-single.lua in NumPy float32."""
    if rule == "rmsprop":
        d, lr, mo, eps = f32(hp["decay"]), hp["lr"], f32(hp["momentum"]), f32(hp["epsilon"])
        st["a"] = st["a"] * d + g * f32(1 - hp["decay"])
        st["s"] = st["s"] * d + (g * g) * f32(1 - hp["decay"])
        rms = np.sqrt(st["s"] - st["a"] * st["a"] + eps)
        st["u"] = st["u"] * mo + (g / rms) * f32(-lr)
        return p + st["u"]
    if rule == "adam":
        b1, b2 = hp["beta1"], hp["beta2"]
        k = (t // hp["step_div"] + 1) if hp.get("step_div") else t
        st["m"] = st["m"] * f32(b1) + g * f32(1 - b1)
        st["v"] = st["v"] * f32(b2) + (g * g) * f32(1 - b2)
        den = np.sqrt(st["v"]) + f32(hp["epsilon"])
        lr_t = hp["lr"] * math.sqrt(1 - b2 ** k) / (1 - b1 ** k)
        return p + (st["m"] / den) * f32(-lr_t)
    if rule == "adamax":
        b1, b2 = hp["beta1"], hp["beta2"]
        st["m"] = st["m"] * f32(b1) + g * f32(1 - b1)
        st["u"] = np.maximum(st["u"] * f32(b2), np.abs(g) + f32(hp["epsilon"]))
        return p + (st["m"] / st["u"]) * f32(-(hp["lr"] / (1 - b1 ** t)))
    if rule == "adagrad":
        st["v"] = st["v"] + g * g
        std = np.sqrt(st["v"]) + f32(hp["epsilon"])
        return p + (g / std) * f32(-(hp["lr"] / (1 + (t - 1) * hp["lrd"])))
    if rule == "adadelta":
        rho = hp["rho"]
        st["v"] = st["v"] * f32(rho) + (g * g) * f32(1 - rho)
        std = np.sqrt(st["v"] + f32(hp["epsilon"]))
        delta = np.sqrt(st["acc"] + f32(hp["epsilon"])) / std * g
        st["acc"] = st["acc"] * f32(rho) + (delta * delta) * f32(1 - rho)
        return p + delta * f32(-hp["lr"])
    raise ValueError(rule)


HP = {
    "rmsprop": dict(decay=0.95, lr=1e-3, momentum=0.9, epsilon=1e-4),
    "adam": dict(beta1=0.9, beta2=0.999, lr=1e-3, epsilon=1e-8, step_div=72),
    "adamax": dict(beta1=0.9, beta2=0.999, lr=2e-3, epsilon=1e-8),
    "adagrad": dict(lr=1e-2, lrd=1e-3, epsilon=1e-10),
    "adadelta": dict(rho=0.9, epsilon=1e-6, lr=1.0),
}


@pytest.mark.parametrize("rule", list(HP))
def test_single_worker_rules_push_parameters(rule):
    gs = _grads(6, seed=3)
    w0 = np.linspace(-1, 1, 257).astype(f32)
    w = torch.from_numpy(w0.copy())
    ps = LocalPS(torch.zeros(257))
    ps.reset(w, None)
    hp = dict(HP[rule])
    hp.pop("step_div", None)          # the single variants use the textbook exponent (optim-adam-single.lua:28-29)
    cfg = dict(hp, pclient=ps)
    it = iter(gs)
    fn = R.OPTIMIZERS[rule + "single"]
    for _ in gs:
        fn(lambda x: (0.0, torch.from_numpy(next(it).copy())), w, cfg)
    st = {k: np.zeros(257, f32) for k in ("a", "s", "u", "m", "v", "acc")}
    p = w0.copy()
    for t, g in enumerate(gs, start=1):
        p = _np_rule(rule, p, g, st, hp, t)
    assert np.allclose(w.numpy(), p, rtol=1e-6, atol=1e-7)
    assert ps.sent_params == 6 and np.array_equal(ps.center.numpy(), w.numpy())


@pytest.mark.parametrize("rule", list(HP))
def test_server_rules_match_numpy(rule):
    gs = _grads(5, seed=4)
    p0 = np.linspace(-1, 1, 257).astype(f32)
    p = torch.from_numpy(p0.copy())
    st_t = R.rule_state(rule, p)
    for g in gs:
        R.apply_rule(rule, p, torch.from_numpy(g.copy()), st_t, HP[rule])
    st = {k: np.zeros(257, f32) for k in ("a", "s", "u", "m", "v", "acc")}
    q = p0.copy()
    for t, g in enumerate(gs, start=1):
        q = _np_rule(rule, q, g, st, HP[rule], t)
    assert np.allclose(p.numpy(), q, rtol=1e-6, atol=1e-7)


def test_parallel_rule_client_accumulates_raw_gradients():
    """su > 1: the client sums raw gradients and ships them every su-th step (BiCNN/optim-adam.lua:27-43)."""
    class Spy(LocalPS):
        def __init__(self):
            super().__init__(torch.zeros(4))
            self.pushed = []

        def async_send_grad(self):
            self.pushed.append(self.g.clone())

        def async_recv_param(self):
            pass

    ps = Spy()
    w = torch.zeros(4)
    cfg = {"pclient": ps, "su": 3, "mode": "global"}
    for t in range(7):
        R.adam(lambda x: (0.0, torch.full((4,), float(t + 1))), w, cfg)
    assert [float(g[0]) for g in ps.pushed] == [1.0, 2.0 + 3.0 + 4.0, 5.0 + 6.0 + 7.0]
