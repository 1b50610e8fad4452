"""GPU numerics of the exchange kernels with every 'rank' aliased onto one device (shard arithmetic,
put/pull/push/elastic and the fused EAMSGD boundary kernel vs the Lua restatement)."""
import pytest
import torch

from mpit_b200.optim import reference as R
from mpit_b200.parallel.center import ShardedCenter, partition, shard_capacity
from mpit_b200.parallel.symm import LocalPeerHeap

pytestmark = pytest.mark.gpu


def _centers(n, S):
    heap = LocalPeerHeap(S, devices=[0] * S)
    bufs = heap.alloc(shard_capacity(n, S))
    return [ShardedCenter(b, n) for b in bufs]


def _center_as_flat(cs, n):
    parts = partition(n, len(cs))
    return torch.cat([cs[s].buf.local[:sz] for s, (off, sz) in enumerate(parts)])


@pytest.mark.parametrize("n,S", [(351946, 8), (1003, 3), (5, 8), (4096, 1), (43993 * 8 + 5, 8),
                                 (1 << 20, 8), ((1 << 20) + 4096 * 3, 3), (3 << 20, 2)])   # last three: bulk-engine path
def test_put_pull_push_roundtrip(n, S):
    cs = _centers(n, S)
    x = torch.randn(n, device="cuda")
    cs[0].put(x)
    torch.cuda.synchronize()
    assert torch.equal(_center_as_flat(cs, n), x)
    y = torch.empty_like(x)
    cs[S - 1].pull(y)
    assert torch.equal(y, x)
    g = torch.randn(n, device="cuda")
    g0 = g.clone()
    cs[min(1, S - 1)].push_add(g, zero_src=True)
    torch.cuda.synchronize()
    assert torch.equal(_center_as_flat(cs, n), x + g0)
    assert float(g.abs().sum()) == 0.0


def test_push_with_clip_and_scale():
    n, S = 1003, 3
    cs = _centers(n, S)
    p = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda") * 5
    cs[0].push_add(g, p, scale=0.5, l2=0.1, clip=1.0)
    ref = (g + p * 0.1).clamp(-1, 1) * 0.5
    assert torch.allclose(_center_as_flat(cs, n), ref, atol=1e-6)


def test_elastic_exchange_conserves():
    n, S = 1003, 4
    cs = _centers(n, S)
    c0 = torch.randn(n, device="cuda")
    cs[0].put(c0)
    x = torch.randn(n, device="cuda")
    d = torch.empty_like(x)
    cs[2].elastic(x, d, 0.25)
    torch.cuda.synchronize()
    ref_d = (x - c0) * 0.25
    assert torch.equal(d, ref_d)
    assert torch.equal(_center_as_flat(cs, n), c0 + ref_d)


class _LocalPS:
    """In-process stand-in for pclient/pserver: one worker, plain-add server."""

    def __init__(self, center):
        self.center = center
        self.p = self.g = None

    def reset(self, p, g):
        self.p, self.g = p, g

    def async_recv_param(self):
        self.p.copy_(self.center)

    def async_send_grad(self):
        self.center.add_(self.g)

    def ping(self):
        pass

    def wait(self):
        pass


@pytest.mark.parametrize("su,mom", [(1, 0.9), (3, 0.99), (2, 0.0)])
def test_fused_eamsgd_step_matches_reference_bitwise(su, mom):
    n, S, K = 10007, 8, 9
    gen = torch.Generator().manual_seed(5)
    grads = [torch.randn(n, generator=gen) * 0.1 for _ in range(K)]
    w0 = torch.randn(n, generator=gen)
    # --- oracle: asyncsgd/optim-eamsgd.lua restated, single worker
    w = w0.clone()
    ps = _LocalPS(w0.clone())
    cfg = {"lr": 0.05, "mom": mom, "mva": 0.3, "su": su, "pclient": ps}
    it = iter(grads)
    for _ in range(K):
        R.eamsgd(lambda w_: (0.0, next(it).clone()), w, cfg)
    # --- device: fused boundary kernel
    cs = _centers(n, S)
    c = cs[3]
    x = w0.cuda()
    c.put(x)
    v = torch.zeros_like(x) if mom > 0 else None
    d = torch.zeros_like(x)
    g_prev = None
    prev_sync = False
    for t in range(K + 1):
        do_apply = t > 0
        last = t == K
        do_sync = (not last) and t % su == 0
        c.eamsgd_step(x, v, g_prev, None, d, None, None, alpha=0.3, clr=0.05, l2wd=0.0, mom_next=mom,
                      do_apply=do_apply, use_mom=mom > 0, do_lookahead=not last, do_sync=do_sync,
                      has_d_in=do_apply and prev_sync)
        if not last:
            g_prev = grads[t].cuda()
        prev_sync = do_sync
    torch.cuda.synchronize()
    assert torch.equal(x.cpu(), w)
    assert torch.equal(_center_as_flat(cs, n).cpu(), ps.center)


@pytest.mark.parametrize("n,S", [(1 << 20, 8), (3 * 4096 * 100, 3)])
def test_bulk_elastic_and_scaled_push(n, S):
    cs = _centers(n, S)
    assert cs[0].bulk
    c0 = torch.randn(n, device="cuda")
    cs[0].put(c0)
    x = torch.randn(n, device="cuda")
    d = torch.empty_like(x)
    cs[1].elastic(x, d, 0.25)
    torch.cuda.synchronize()
    ref_d = (x - c0) * 0.25
    assert torch.equal(d, ref_d)
    assert torch.equal(_center_as_flat(cs, n), c0 + ref_d)
    g = torch.randn(n, device="cuda")
    g0 = g.clone()
    cs[S - 1].push_add(g, scale=0.5, zero_src=True)
    torch.cuda.synchronize()
    assert torch.allclose(_center_as_flat(cs, n), c0 + ref_d + g0 * 0.5, atol=1e-6)
    assert float(g.abs().sum()) == 0.0


def test_device_window_put_get_accumulate_single_gpu():
    """MPI_Put / MPI_Get / MPI_Accumulate on symmetric memory (mpifuncs.c:1656,1131,9), ranks aliased onto one GPU,
    odd displacements and lengths (scalar tail path) included.  No fence here: barrier kernels of different ranks
    must not share one GPU."""
    from mpit_b200.parallel.window import DeviceWindow
    S, n = 3, 1000
    heap = LocalPeerHeap(S, devices=[0] * S)
    bufs, sigs = heap.alloc(n), heap.signal_pads()
    wins = [DeviceWindow(None, n, buf=bufs[r], sig=sigs[r]) for r in range(S)]
    a = torch.randn(257, device="cuda")
    wins[0].put(a, target_rank=2, target_disp=13)
    wins[1].accumulate(a, target_rank=2, target_disp=13, scale=0.5)
    wins[0].unlock(2)
    want = torch.zeros(n, device="cuda")
    want[13:13 + 257] = a * 1.5
    assert torch.allclose(wins[2].local, want)
    out = torch.empty(100, device="cuda")
    wins[1].get(out, target_rank=2, target_disp=16, count=100)      # 16-byte aligned: vector path
    wins[1].flush(2)
    assert torch.equal(out, wins[2].local[16:116])
    with pytest.raises(ValueError):
        wins[0].put(a, target_rank=1, target_disp=n - 10)
    with pytest.raises(ValueError):
        wins[0].get(out, target_rank=5)


def test_centre_checkpoint_roundtrip(tmp_path):
    """ShardedCenter.save: every owner writes its (device) shard into one flat Torch7 file; load restores it."""
    from mpit_b200 import t7
    n, S = 351946, 3
    cs = _centers(n, S)
    theta = torch.randn(n, device="cuda")
    cs[0].put(theta)
    torch.cuda.synchronize()
    path = str(tmp_path / "centre.t7")
    for c in cs:                                       # the aliased "ranks" take their turns; rank 0 creates the file
        c.save(path, barrier=lambda: None)
    assert torch.equal(t7.load(path), theta.cpu())
    cs[0].put(torch.zeros(n, device="cuda"))
    torch.cuda.synchronize()
    for c in cs:
        c.load(path)
    assert torch.equal(_center_as_flat(cs, n), theta)


@pytest.mark.parametrize("su,l2wd", [(1, 0.0), (3, 0.0), (2, 1e-3)])
def test_downpour_fused_boundary_matches_reference_bitwise(su, l2wd):
    """One worker against a 3-shard centre: the single boundary kernel (accumulate, returning NVLink accumulate,
    continue from the owner's value) reproduces optim-downpour.lua:24-54 driven through a one-server restatement
    (push p += g', then pull) bit for bit."""
    from mpit_b200.optim.device import DeviceDownpour
    n, S, steps, lr = 10007, 3, 7, 0.05
    cs = _centers(n, S)
    grads = [torch.randn(n, generator=torch.Generator().manual_seed(50 + t)) * 0.1 for t in range(steps)]

    class Eng:
        def __init__(self):
            self.wpk = torch.zeros(n, device="cuda")
            self.gpk = torch.zeros(n, device="cuda")
            self.gmap = None
            self.wmap = None
            self.t = 0

        def load_flat(self, x):
            self.wpk.copy_(x)

        def fwd_bwd(self):
            self.gpk.copy_(grads[self.t])
            self.t += 1

    theta0 = torch.linspace(-1, 1, n)
    opt = DeviceDownpour(Eng(), theta0, cs[1], lr=lr, su=su, l2wd=l2wd, seed_center=True, use_graphs=False)
    # restatement (reference.py downpour with a server that adds and returns the parameters)
    w, centre, acc = theta0.clone(), theta0.clone(), torch.zeros(n)
    f32 = lambda v: torch.tensor(v, dtype=torch.float32)
    for t in range(steps):
        opt.step()
        g = grads[t].clone()
        if l2wd != 0:
            g = g + w * f32(l2wd)
        g = g * f32(-lr)
        acc = acc + g
        if t % su == 0:
            centre = centre + acc
            w = centre.clone()
            acc.zero_()
        else:
            w = w + g
    torch.cuda.synchronize()
    assert torch.equal(opt.x.cpu(), w)
    assert torch.equal(_center_as_flat(cs, n).cpu(), centre)
    assert torch.equal(opt.engine.wpk.cpu(), w), "operand copy refreshed in the same pass"
    assert torch.equal(opt.acc.cpu(), acc)
