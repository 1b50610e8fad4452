"""GPU tests of the asynchronous parameter server with server-side rules (csrc/kernels/ps_async.cu), every 'rank'
aliased onto one device so that the one-GPU tier covers the multi-client logic: arrival-order application against the
restatement of /root/reference/BiCNN/pserver.lua:115-205, partial arrivals (an owner never waits), flow control, the
su-step accumulation / L1 / L2 / clamp prologue of the clients (/root/reference/BiCNN/optim-adam.lua:24-57,
bicnn.lua:398-409), the native server thread and CUDA-graph capture of a whole step."""
import time

import pytest
import torch

from mpit_b200.optim import reference as R
from mpit_b200.parallel.center import partition
from mpit_b200.parallel.ps_device import AsyncDevicePS
from mpit_b200.parallel.symm import LocalPeerHeap

pytestmark = pytest.mark.gpu

HYPER = {
    "rmsprop": dict(decay=0.95, lr=1e-3, momentum=0.9, epsilon=1e-4),
    "adam": dict(beta1=0.9, beta2=0.999, lr=1e-3, epsilon=1e-8, step_div=3),
    "adamax": dict(beta1=0.9, beta2=0.999, lr=2e-3, epsilon=1e-8),
    "adagrad": dict(lr=1e-2, lrd=1e-3, epsilon=1e-10),
    "adadelta": dict(rho=0.9, epsilon=1e-6, lr=1.0),
    "add": dict(),
}


def _grad(n, r, t):
    return torch.randn(n, generator=torch.Generator().manual_seed(1000 * r + t)) * 0.1


def _todo_order(ps, k):
    return ps.ctl.local[64:64 + k].tolist()


@pytest.mark.parametrize("rule", list(HYPER))
def test_rules_applied_in_arrival_order(rule):
    n, S, rounds = 10007, 3, 4
    group = AsyncDevicePS.local_group(LocalPeerHeap(S, devices=[0] * S), n, rule, HYPER[rule])
    theta0 = torch.linspace(-1, 1, n)
    group[0].center.put(theta0.cuda())
    parts = partition(n, S)
    oracle = [theta0[o:o + sz].clone() for o, sz in parts]
    state = [R.rule_state(rule, p) for p in oracle]
    orders = [[2, 0, 1], [1, 2, 0], [0, 1, 2], [2, 1, 0]]
    for t in range(rounds):
        for c in orders[t]:                       # one stream: arrival order == launch order
            group[c].push(gflat=_grad(n, c, t).cuda())
        for s in range(S):
            group[s].drain()
        torch.cuda.synchronize()
        for s, (o, sz) in enumerate(parts):
            assert _todo_order(group[s], S) == orders[t]
            for c in orders[t]:
                R.apply_rule(rule, oracle[s], _grad(n, c, t)[o:o + sz], state[s], HYPER[rule])
    for s, (o, sz) in enumerate(parts):
        got = group[s].center.buf.local[:sz].cpu()
        assert torch.allclose(got, oracle[s], rtol=1e-6, atol=1e-7), (rule, s, float((got - oracle[s]).abs().max()))
        assert group[s].applied == rounds * S and group[s].arrivals == rounds * S
        group[s].check_timeout()
    x = torch.empty(n, device="cuda")
    group[1].pull(x)
    assert torch.allclose(x.cpu(), torch.cat(oracle), rtol=1e-6, atol=1e-7)


def test_owner_applies_what_has_arrived_and_never_waits():
    n, S = 4099, 4
    group = AsyncDevicePS.local_group(LocalPeerHeap(S, devices=[0] * S), n, "add", {})
    one = torch.ones(n, device="cuda")
    group[1].push(gflat=one)
    group[3].push(gflat=one * 2)
    for s in range(S):
        group[s].drain()                       # clients 0 and 2 have not pushed: nothing waits for them
    torch.cuda.synchronize()
    x = torch.empty(n, device="cuda")
    group[0].pull(x)
    assert float(x.min()) == 3.0 and float(x.max()) == 3.0
    assert all(g.applied == 2 for g in group)
    for s in range(S):
        group[s].drain()                       # an empty drain is a no-op
    group[0].push(gflat=one * 4)
    group[1].push(gflat=one * 8)               # second push of client 1: its slots were acknowledged above
    for s in range(S):
        group[s].drain(max_take=1)             # one gradient per drain ...
    group[2].pull(x)
    torch.cuda.synchronize()
    assert float(x.min()) == 7.0 and float(x.max()) == 7.0
    for s in range(S):
        group[s].drain(max_take=1)             # ... the other one on the next
    group[2].pull(x)
    torch.cuda.synchronize()
    assert float(x.min()) == 15.0 and float(x.max()) == 15.0
    for g in group:
        assert g.applied == 4
        g.check_timeout()
    assert [g.pushes for g in group] == [1, 2, 0, 1]


def test_client_prologue_accumulation_reg_clip_and_maps():
    """su = 2: the client sums two raw gradients (each regularised and clamped as in bicnn.lua:398-409) and pushes the
    sum; the dense part of the gradient comes from a packed buffer through an index table."""
    n, S, base = 5003, 2, 1000
    group = AsyncDevicePS.local_group(LocalPeerHeap(S, devices=[0] * S), n, "add", {})
    perm = torch.randperm(n - base)
    gmap = perm.to(torch.int32).cuda()
    x = torch.randn(n, device="cuda")
    acc = torch.zeros(n, device="cuda")
    active = torch.ones(1, dtype=torch.int32, device="cuda")
    reg = torch.zeros(2, device="cuda")
    want = torch.zeros(n, device="cuda")
    for t in range(2):
        flat = torch.randn(n, device="cuda") * 3
        gpk = torch.randn(n - base, device="cuda") * 3
        full = torch.cat([flat[:base], gpk[perm.cuda()]])
        want += (full + torch.sign(x) * 0.01 + x * 0.1).clamp(-1.0, 1.0)
        group[0].push(gflat=flat, gpk=gpk, gmap=gmap, map_base=base, x=x, acc=acc, active=active, reg_loss=reg,
                      l1=0.01, l2=0.1, clip=1.0, do_push=(t == 1))
        if t == 0:
            assert torch.allclose(acc, want, atol=1e-6) and group[0].pushes == 0
    for s in range(S):
        group[s].drain()
    y = torch.empty(n, device="cuda")
    group[1].pull(y)
    torch.cuda.synchronize()
    assert torch.allclose(y, want, atol=1e-5), float((y - want).abs().max())
    assert float(acc.abs().max()) == 0.0
    ref_reg = torch.stack([0.01 * x.abs().sum(), 0.05 * (x * x).sum()]) * 2
    assert torch.allclose(reg, ref_reg, rtol=1e-4)
    # an inactive step (no margin violator: zero gradient, bicnn.lua:361-374) skips regularisation and clamping
    active.zero_()
    group[0].push(gflat=torch.zeros(n, device="cuda"), x=x, active=active, l1=0.01, l2=0.1, clip=1.0)
    for s in range(S):
        group[s].drain()
    group[1].pull(y)
    torch.cuda.synchronize()
    assert torch.allclose(y, want, atol=1e-5)


class _FakeEngine:
    """Gradient source with the engine interface; the gradient of step t is a fixed function of t."""

    def __init__(self, n, steps):
        self.n = n
        self.table = torch.stack([_grad(n, 7, t) for t in range(steps)]).cuda()
        self.tdev = torch.zeros(1, dtype=torch.int64, device="cuda")
        self.wpk = torch.zeros(n, device="cuda")
        self.gpk = torch.zeros(n, device="cuda")
        self.gmap = None
        self.wmap = None

    def load_flat(self, x):
        self.wpk.copy_(x)

    def fwd_bwd(self):                         # capturable: everything on the device
        self.gpk.copy_(self.table.index_select(0, self.tdev).view(-1))
        self.tdev.add_(1)


@pytest.mark.parametrize("rule,su,graphs", [("adam", 1, True), ("adagrad", 2, True), ("rmsprop", 1, False),
                                            ("adamax", 3, True), ("adadelta", 1, True)])
def test_one_client_trajectory_equals_reference(rule, su, graphs):
    """One client, inline server, the whole step in a CUDA graph: push -> drain -> pull reproduces the reference's
    client + server pair update for update (optim-adam.lua:24-57 + pserver.lua:115-205)."""
    from mpit_b200.optim.device import DeviceRuleParallel
    from mpit_b200.parallel.center import ShardedCenter
    n, steps = 6151, 9
    heap = LocalPeerHeap(1, devices=[0])

    class _Heap:                                # SymmetricHeap interface over the single local rank
        nranks, rank = 1, 0
        def alloc(self, numel, dtype=torch.float32, multicast=False):
            return heap.alloc(numel, dtype)[0]

    h = _Heap()
    center = ShardedCenter.create(h, n)
    eng = _FakeEngine(n, steps)
    theta0 = torch.linspace(-1, 1, n)
    opt = DeviceRuleParallel(eng, theta0, rule, HYPER[rule], center, h, su=su, clip=0.15, use_graphs=graphs)
    p = theta0.clone()
    st = R.rule_state(rule, p)
    acc = torch.zeros(n)
    for t in range(steps):
        opt.step()
        acc += _grad(n, 7, t).clamp(-0.15, 0.15)
        if t % su == 0:
            R.apply_rule(rule, p, acc if su > 1 else _grad(n, 7, t).clamp(-0.15, 0.15), st, HYPER[rule])
            acc.zero_()
        torch.cuda.synchronize()
        assert torch.equal(eng.wpk, opt.x), "operand copy refreshed by the pull"
    got = opt.x.cpu()
    assert torch.allclose(got, p, rtol=1e-6, atol=1e-7), float((got - p).abs().max())
    assert opt.ps.applied == (steps + su - 1) // su
    opt.ps.check_timeout()


def test_native_server_thread_drains_without_the_training_stream():
    n, S = 20011, 2
    group = AsyncDevicePS.local_group(LocalPeerHeap(S, devices=[0] * S), n, "adam", HYPER["adam"])
    theta0 = torch.zeros(n, device="cuda")
    group[0].center.put(theta0)
    torch.cuda.synchronize()
    for g in group:
        g.start_server(interval_us=50)
    pushes = 6
    for t in range(pushes):
        for c in range(S):
            group[c].push(gflat=_grad(n, c, t).cuda())       # flow control waits for the thread's acknowledgements
    torch.cuda.current_stream().synchronize()
    deadline = time.time() + 10
    while any(g.applied < pushes * S for g in group) and time.time() < deadline:
        time.sleep(0.01)
    rounds = [g.stop_server() for g in group]
    assert all(g.applied == pushes * S for g in group), [g.applied for g in group]
    assert all(r >= 1 for r in rounds)
    x = torch.empty(n, device="cuda")
    group[0].pull(x)
    torch.cuda.synchronize()
    assert torch.isfinite(x).all() and float(x.abs().max()) > 0
    for g in group:
        g.check_timeout()
