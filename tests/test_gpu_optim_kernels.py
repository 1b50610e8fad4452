"""GPU numerics: fused optimizer kernels vs the torch restatements of the Lua rules."""
import pytest
import torch

from mpit_b200 import ops
from mpit_b200.optim import reference as R

pytestmark = pytest.mark.gpu
N = 100_003   # odd: exercises the scalar tail


def _grads(k, n=N, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(n, generator=g) * 0.1 for _ in range(k)]


@pytest.mark.parametrize("mom,l2wd", [(0.0, 0.0), (0.9, 0.0), (0.99, 1e-3)])
def test_local_step_matches_msgd_bitwise(mom, l2wd):
    gs = _grads(6)
    w0 = torch.randn(N, generator=torch.Generator().manual_seed(1))
    # oracle
    w = w0.clone()
    cfg = {"lr": 0.05, "mom": mom, "l2wd": l2wd}
    it = iter(gs)
    for _ in gs:
        R.msgd(lambda w_: (0.0, next(it).clone()), w, cfg)
    # device: lookahead / apply halves
    x = w0.cuda()
    v = torch.zeros_like(x) if mom > 0 else None
    for g in gs:
        if mom > 0:
            ops.local_step(x, v, None, None, None, None, None, mom_next=mom, do_apply=False, use_mom=True,
                           do_lookahead=True)
        ops.local_step(x, v, g.cuda(), None, None, None, None, clr=0.05, l2wd=l2wd, do_apply=True,
                       use_mom=mom > 0, do_lookahead=False)
    torch.cuda.synchronize()
    assert torch.equal(x.cpu(), w)
    if mom > 0:
        assert torch.equal(v.cpu(), cfg["vt"])


def test_local_step_with_maps_and_pack():
    n = 1000
    perm = torch.randperm(n)
    wmap = perm.to(torch.int32).cuda()
    x = torch.randn(n).cuda()
    x0 = x.clone()
    gpk = torch.randn(n).cuda()
    wpk = torch.zeros(n).cuda()
    ops.local_step(x, None, gpk, wmap, None, wpk, wmap, clr=0.1, do_apply=True)
    ref = x0 + gpk[perm.cuda()] * torch.tensor(-0.1).float()
    assert torch.equal(x, ref)
    assert torch.equal(wpk[perm.cuda()], x)


HYPER = {
    "rmsprop": dict(decay=0.95, lr=1e-3, momentum=0.9, epsilon=1e-4),
    "adam": dict(beta1=0.9, beta2=0.999, lr=1e-3, epsilon=1e-8),
    "adamax": dict(beta1=0.9, beta2=0.999, lr=2e-3, epsilon=1e-8),
    "adagrad": dict(lr=1e-2, lrd=1e-3, epsilon=1e-10),
    "adadelta": dict(rho=0.9, epsilon=1e-6, lr=1.0),
    "add": dict(),
}


@pytest.mark.parametrize("rule", list(HYPER))
def test_rule_step_matches_reference(rule):
    hp = HYPER[rule]
    gs = _grads(5, seed=3)
    p0 = torch.randn(N, generator=torch.Generator().manual_seed(2))
    p = p0.clone()
    st = R.rule_state(rule, p)
    for g in gs:
        R.apply_rule(rule, p, g, st, hp)
    x = p0.cuda()
    z = lambda: torch.zeros_like(x)
    s0, s1, s2 = z(), z(), z()
    for t, g in enumerate(gs, start=1):
        ops.rule_step(rule, x, g.cuda(), ops.rule_hyper(rule, t, **hp), s0, s1, s2)
    torch.cuda.synchronize()
    got = x.cpu()
    # same op order and roundings; sqrt/div are IEEE on both sides
    assert torch.allclose(got, p, rtol=1e-6, atol=1e-7), float((got - p).abs().max())


def test_adam_server_step_div_quirk():
    hp = dict(HYPER["adam"], step_div=3)
    h5 = ops.rule_hyper("adam", 5, **hp)   # k = floor(5/3)+1 = 2
    h2 = ops.rule_hyper("adam", 2, **HYPER["adam"])
    assert h5 == h2


def test_downpour_accum():
    n = 4099
    x = torch.randn(n).cuda(); acc = torch.zeros(n).cuda(); g = torch.randn(n).cuda()
    x0 = x.clone()
    ops.downpour_accum(x, acc, g, lr=0.1, move_local=True)
    step = g * torch.tensor(-0.1).float()
    assert torch.equal(acc, step) and torch.equal(x, x0 + step)
    ops.downpour_accum(x, acc, g, lr=0.1, move_local=False)
    assert torch.equal(acc, step + step) and torch.equal(x, x0 + step)


def test_grad_reg_clip():
    n = 1000
    g = torch.randn(n).cuda() * 3; p = torch.randn(n).cuda()
    ref = (g + torch.sign(p) * 0.01 + p * 0.1).clamp(-1.0, 1.0)
    ops.grad_reg_clip(g, p, l1=0.01, l2=0.1, clip=1.0)
    assert torch.allclose(g, ref, atol=1e-6)
