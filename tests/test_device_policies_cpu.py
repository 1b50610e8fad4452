"""Host-side pieces of the device paths that need no GPU: the scalar preparation the rule kernels are launched with
(``ops.optim_ops.rule_hyper``: step sizes, the server's ``floor(t / stepDivAdam) + 1`` exponent) against the torch
restatement of /root/reference/BiCNN/pserver.lua:123-195, the option table -> rule mapping of the device parameter
server against the host server's, and the tiling / split-K policies of the pull-fused GEMM."""
import numpy as np
import pytest
import torch

from mpit_b200.optim import reference as R
from mpit_b200.ops.optim_ops import rule_hyper

HYPER = {
    "rmsprop": dict(decay=0.95, lr=0.002, momentum=0.9, epsilon=1e-4),
    "adam": dict(lr=0.01, beta1=0.9, beta2=0.999, epsilon=1e-8, step_div=3),
    "adam0": dict(lr=0.01, beta1=0.9, beta2=0.999, epsilon=1e-8, step_div=0),
    "adamax": dict(lr=0.02, beta1=0.9, beta2=0.999, epsilon=1e-8),
    "adagrad": dict(lr=0.05, lrd=0.01, epsilon=1e-10),
    "adadelta": dict(rho=0.9, epsilon=1e-6, lr=1.0),
}


def _kernel_element(rule, p, g, s, h):
    """csrc/kernels/optim_math.cuh:rule_apply, one element, in float32 like the kernel (one rounding per operation)."""
    f = np.float32
    a, b, c, d, eps, step = (f(x) for x in h)
    if rule == "rmsprop":
        s[0] = f(f(s[0] * a) + f(b * g))
        s[1] = f(f(s[1] * a) + f(b * f(g * g)))
        rms = np.sqrt(f(f(s[1] - f(s[0] * s[0])) + eps))
        s[2] = f(f(s[2] * c) + f(step * f(g / rms)))
        return f(p + s[2])
    if rule == "adam":
        s[0] = f(f(s[0] * a) + f(b * g))
        s[1] = f(f(s[1] * c) + f(d * f(g * g)))
        return f(p + f(step * f(s[0] / f(np.sqrt(s[1]) + eps))))
    if rule == "adamax":
        s[0] = f(f(s[0] * a) + f(b * g))
        s[1] = max(f(s[1] * c), f(abs(g) + eps))
        return f(p + f(step * f(s[0] / s[1])))
    if rule == "adagrad":
        s[0] = f(s[0] + f(g * g))
        return f(p + f(step * f(g / f(np.sqrt(s[0]) + eps))))
    if rule == "adadelta":
        s[0] = f(f(s[0] * a) + f(b * f(g * g)))
        std = np.sqrt(f(s[0] + eps))
        delta = f(f(np.sqrt(f(s[1] + eps)) / std) * g)
        p = f(p + f(step * delta))
        s[1] = f(f(s[1] * a) + f(b * f(delta * delta)))
        return p
    raise ValueError(rule)


@pytest.mark.parametrize("name", sorted(HYPER))
def test_kernel_scalars_reproduce_the_reference_rule(name):
    rule = name.rstrip("0")
    hp = HYPER[name]
    gen = torch.Generator().manual_seed(7)
    n, steps = 64, 9
    p_ref = torch.randn(n, generator=gen)
    st = R.rule_state(rule, p_ref)
    p = p_ref.numpy().copy()
    s = [np.zeros(n, np.float32) for _ in range(3)]
    for t in range(1, steps + 1):
        g = torch.randn(n, generator=gen) * 0.3
        R.apply_rule(rule, p_ref, g, st, hp)
        h = rule_hyper(rule, t, **hp)                              # what the kernel launch of update t is given
        for i in range(n):
            si = [x[i] for x in s]
            p[i] = _kernel_element(rule, np.float32(p[i]), np.float32(g[i]), si, h)
            for x, v in zip(s, si):
                x[i] = v
        assert st["t"] == t
        err = np.abs(p - p_ref.numpy()).max()
        assert err <= 2e-6 * max(1.0, float(p_ref.abs().max())), (name, t, err)


def test_adam_exponent_follows_the_servers_step_div_rule():
    # pserver.lua:151-152: the bias-correction exponent is floor(t / stepDivAdam) + 1, not t
    hp = dict(lr=0.01, beta1=0.9, beta2=0.999, epsilon=1e-8)
    for t in range(1, 12):
        k = t // 3 + 1
        want = -0.01 * np.sqrt(1 - 0.999 ** k) / (1 - 0.9 ** k)
        assert abs(rule_hyper("adam", t, step_div=3, **hp)[5] - want) < 1e-12
        want0 = -0.01 * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        assert abs(rule_hyper("adam", t, step_div=0, **hp)[5] - want0) < 1e-12


def test_device_server_gets_the_same_rule_and_hyper_parameters_as_the_host_server():
    from mpit_b200.launch.plaunch import parse
    from mpit_b200.ps import server_rule_config
    from mpit_b200.workloads.bicnn import device_rule_for
    for name in ("rmsprop", "adam", "adamax", "adagrad", "adadelta"):
        opt = parse(["-optimization", name])
        rule, hyper, gscale = device_rule_for(opt)
        rule_h, hyper_h = server_rule_config(opt)
        assert rule == rule_h == name and gscale == 1.0
        for k, v in hyper.items():
            assert hyper_h[k] == v, (name, k)
    opt = parse(["-optimization", "downpour", "-learningRate", "0.05"])
    assert device_rule_for(opt) == ("add", {}, -0.05)              # the client pushes -lr * (accumulated g)
    with pytest.raises(ValueError):
        device_rule_for(parse(["-optimization", "rmsprop", "-modeRMSProp", "local"]))
    with pytest.raises(ValueError):
        device_rule_for(parse(["-optimization", "sgd"]))


def test_pulled_gemm_tiling_policies():
    from mpit_b200.ops.pulled_linear import _split_k, _tiles_per_cta
    for m_tiles in range(1, 17):
        for col_tiles in (2, 4, 16, 32, 64, 128):
            mt = _tiles_per_cta(m_tiles, col_tiles)
            assert mt in (1, 2, 4) and mt <= max(1, m_tiles)
            if mt > 1:                                            # reuse is only taken when enough CTAs remain
                assert -(-m_tiles // mt) * col_tiles >= 64
    for ctas in (1, 4, 16, 32, 64, 100, 148, 300):
        for num_kb in (2, 8, 16, 64, 128, 129, 512):
            ks = _split_k(ctas, num_kb)
            assert ks in (1, 2, 4, 8) and num_kb % ks == 0
            if ks > 1:
                assert ctas * ks <= 148 and num_kb // ks >= 8     # one wave, and slices long enough to pipeline
    assert _split_k(16, 128) == 8 and _split_k(32, 128) == 4 and _split_k(64, 128) == 2 and _split_k(128, 128) == 1
