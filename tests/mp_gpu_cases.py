"""Multi-GPU test bodies: run with ``python -m torch.distributed.run --nproc-per-node N tests/mp_gpu_cases.py <case>``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from mpit_b200.parallel import SymmetricHeap, ShardedCenter, DeviceCollectives
from mpit_b200.parallel.center import partition


def setup():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    return rank, world, dev


def sync(dev):
    torch.cuda.synchronize(dev)
    dist.barrier()
    torch.cuda.synchronize(dev)


def case_center(rank, world, dev, heap):
    n = 351946
    c = ShardedCenter.create(heap, n)
    x0 = torch.arange(n, dtype=torch.float32, device=dev) * 1e-3
    if rank == 0:
        c.put(x0)
    sync(dev)
    y = torch.empty(n, device=dev)
    c.pull(y)
    torch.cuda.synchronize()
    assert torch.equal(y, x0), "pull after seed"
    off, sz = partition(n, world)[rank]
    assert torch.equal(c.local_shard, x0[off:off + sz]), "my shard holds the reference partition"
    sync(dev)
    # conservation of pushes: every rank accumulates (rank+1) everywhere
    g = torch.full((n,), float(rank + 1), device=dev)
    c.push_add(g)
    sync(dev)
    c.pull(y)
    torch.cuda.synchronize()
    tot = world * (world + 1) / 2
    assert torch.allclose(y, x0 + tot), float((y - x0 - tot).abs().max())
    sync(dev)
    # elastic: d_i = alpha (x_i - centre_seen_by_i); centre += d_i.  Ranks race by design
    # ("no lock ... expect inconsistent read", asyncsgd/pserver.lua:81), so each d_i depends on the
    # interleaving -- what must hold exactly is conservation: centre_after = centre_before + sum_i d_i.
    before = y.clone()
    xi = x0 + tot + (rank + 1)
    d = torch.empty(n, device=dev)
    c.elastic(xi, d, 0.1)
    sync(dev)
    lo, hi = 0.1 * ((rank + 1) - 0.1 * tot) - 1e-3, 0.1 * (rank + 1) + 1e-3
    assert float(d.min()) >= lo and float(d.max()) <= hi, (float(d.min()), float(d.max()), lo, hi)
    dsum = d.clone()
    dist.all_reduce(dsum)
    c.pull(y)
    torch.cuda.synchronize()
    assert torch.allclose(y, before + dsum, atol=2e-3), float((y - before - dsum).abs().max())


def case_collectives(rank, world, dev, heap):
    for n, algo in ((6, "oneshot"), (1000, "oneshot"), (1 << 20, "twoshot"), ((1 << 20) + 3, "twoshot"), (4096, "twoshot"),
                    (351946, "oneshot"), (351946, "twoshot"), ((9 << 20) + 1, "twoshot")):
        coll = DeviceCollectives(heap, n)
        x = torch.full((n,), float(rank + 1), device=dev) + torch.arange(n, device=dev) % 7
        out = coll.allreduce(x, algo=algo)
        torch.cuda.synchronize()
        want = world * (world + 1) / 2 + world * (torch.arange(n, device=dev) % 7)
        assert torch.allclose(out, want), (n, algo, float((out - want).abs().max()))
        sync(dev)
        out = coll.allreduce(x, algo=algo, average=True)
        torch.cuda.synchronize()
        assert torch.allclose(out, want / world, atol=1e-5)
        sync(dev)
        # fused elastic pull toward the mean: fresh values, and the output buffer poisoned first so that a tail
        # pass reading a chunk before its owner wrote it cannot pass on leftovers of the previous call
        x2 = x * 0.5 + (torch.arange(n, device=dev) % 5) * (rank + 1)
        coll.out.local.fill_(float("nan"))
        sync(dev)
        xe = x2.clone()
        coll.allreduce(x2, algo=algo, average=True, elastic_x=xe, alpha=0.25)
        torch.cuda.synchronize()
        mean2 = want / world * 0.5 + (torch.arange(n, device=dev) % 5) * ((world + 1) / 2)
        assert torch.allclose(xe, x2 - 0.25 * (x2 - mean2), atol=1e-4), (n, algo, float((xe - (x2 - 0.25 * (x2 - mean2))).abs().max()))
        sync(dev)
        # back-to-back launches without a host sync in between (epoch words, buffer reuse)
        for rep in range(3):
            out = coll.allreduce(x, algo=algo)
        torch.cuda.synchronize()
        assert torch.allclose(out, want), (n, algo, "back to back")
        sync(dev)
        # non-blocking: Test / Wait state machine (test/testireduceall.lua:27-39)
        req = coll.iallreduce(x, algo=algo)
        req.wait(host=True)
        assert req.test()
        assert torch.allclose(coll.out.local[:n], want)
        sync(dev)
        # broadcast from the last rank
        root = world - 1
        coll.out.local[:n].fill_(float(rank))
        sync(dev)
        coll.bcast(n, root)
        torch.cuda.synchronize()
        assert float(coll.out.local[:n].min()) == root and float(coll.out.local[:n].max()) == root
        sync(dev)
        coll.check_timeout()
    coll.barrier()
    torch.cuda.synchronize()
    heap.check_timeout()


def case_eamsgd_train(rank, world, dev, heap):
    """Asynchronous EAMSGD over the sharded centre (mlaunch.lua:50-60: mva = 0.9 / workers) on a set that can be
    learned: every worker's loss has to come down, worker and centre stay finite and close, and the CENTRE -- what
    EASGD returns -- has to have learned too (its test loss is below the ln 10 of a uniform guess, its accuracy above
    chance)."""
    import math
    from mpit_b200.workloads.cifar import CifarTrainer, CifarOpts, learnable_cifar
    from mpit_b200.models.cifar_cnn import random_crop_offsets
    B, steps = 32, 120
    opts = CifarOpts(name="eamsgd", lr=5e-2, mom=0.9, mb=B, su=2, mva=0.9 / world, seed=1 + rank, use_graphs=True)
    tr = CifarTrainer(opts, dev, heap)
    x, y = learnable_cifar(8 * B, seed=rank)
    g = torch.Generator().manual_seed(rank)
    losses = []
    for i in range(steps):
        j = (i % 8) * B
        tr.train_step(x[j:j + B], y[j:j + B], random_crop_offsets(B, g))
        losses.append(tr.last_loss())
    theta = tr.params()
    sync(dev)
    centre = tr.center.gather_full()
    torch.cuda.synchronize()
    assert torch.isfinite(theta).all() and torch.isfinite(centre).all()
    first, last = sum(losses[:10]) / 10, sum(losses[-10:]) / 10
    rel = float((theta - centre).norm() / centre.norm())
    tx, ty = learnable_cifar(4 * B, seed=999)
    tr.opt.x.copy_(centre)
    res = tr.test(tx, ty)
    print(f"[rank {rank}] eamsgd: worker loss {first:.3f} -> {last:.3f}, |theta - centre| / |centre| = {rel:.3f}, "
          f"centre test loss {res['loss']:.3f} accuracy {res['acc']:.3f}", flush=True)
    # measured at 2 GPUs: loss 2.29 -> 0.01 / 0.10, distance 0.01 / 0.04, centre accuracy 1.0
    assert last < 0.6 * first, (first, last)
    assert rel < 0.3, rel
    assert res["loss"] < math.log(10.0) and res["acc"] > 0.3, res
    heap.check_timeout()

PS_HYPER = {
    "adam": dict(lr=0.01, beta1=0.9, beta2=0.999, epsilon=1e-8, step_div=3),
    "adagrad": dict(lr=1e-2, lrd=1e-3, epsilon=1e-10),
    "adadelta": dict(rho=0.9, epsilon=1e-6, lr=1.0),
    "rmsprop": dict(decay=0.95, lr=1e-3, momentum=0.9, epsilon=1e-4),
}


def _ps_grad(n, r, t):
    return torch.randn(n, generator=torch.Generator().manual_seed(1000 * r + t)) * 0.1


def case_server_rules(rank, world, dev, heap):
    """Server-side rules on the asynchronous device path: every owner applies the gradients of a round in the order
    their arrival tickets were drawn (read back from the device) and must end up, on its shard, exactly where the CPU
    restatement of BiCNN/pserver.lua:115-205 ends up when fed the same order."""
    from mpit_b200.optim import reference as R
    from mpit_b200.parallel.ps_device import AsyncDevicePS
    n, rounds = 10007, 4
    for rule, hyper in PS_HYPER.items():
        center = ShardedCenter.create(heap, n)
        ps = AsyncDevicePS(heap, n, rule, hyper, center)
        theta0 = torch.linspace(-1, 1, n)
        if rank == 0:
            center.put(theta0.to(dev))
        sync(dev)
        off, sz = center.parts[rank]
        p = theta0[off:off + sz].clone()
        st = R.rule_state(rule, p)
        for t in range(rounds):
            pushers = list(range(world)) if t != 2 else list(range(0, world, 2))     # round 2: only the even ranks push
            if rank in pushers:
                ps.push(gflat=_ps_grad(n, rank, t).to(dev))
            sync(dev)                      # every gradient of the round has arrived
            ps.drain()
            torch.cuda.synchronize()
            order = ps.ctl.local[64:64 + len(pushers)].tolist()
            assert sorted(order) == pushers, (order, pushers)
            for c in order:
                R.apply_rule(rule, p, _ps_grad(n, c, t)[off:off + sz], st, hyper)
            sync(dev)
        got = center.local_shard.cpu()
        assert torch.allclose(got, p, rtol=1e-6, atol=1e-7), (rule, float((got - p).abs().max()))
        assert ps.applied == (rounds - 1) * world + len(range(0, world, 2))
        x = torch.empty(n, device=dev)
        ps.pull(x)
        torch.cuda.synchronize()
        assert torch.equal(x[off:off + sz].cpu(), got)
        ps.check_timeout()
        sync(dev)


def case_ps_stress(rank, world, dev, heap):
    """Skewed ranks on the asynchronous path (the straggler simulation of /root/reference/BiCNN/ptest2.lua:57-70:
    (rank+1-size/2)^2 busy iterations per round), owners served by their native server threads.  Invariants:
    every push is applied exactly once on every shard (update count = sum of pushes), nothing is lost or doubled
    (plain-add rule: centre = theta0 + sum of everything pushed), no wait times out, and the fast ranks are not
    slowed to the pace of the slowest one."""
    import time
    from mpit_b200.parallel.ps_device import AsyncDevicePS
    n = 3_458_600 if world > 2 else 400_003          # BiCNN-sized vector on the larger boxes
    rounds = 24
    skew = (rank + 1 - world / 2) ** 2
    for rule, hyper in (("add", {}), ("adam", PS_HYPER["adam"])):
        center = ShardedCenter.create(heap, n)
        ps = AsyncDevicePS(heap, n, rule, hyper, center)
        if rank == 0:
            center.put(torch.zeros(n, device=dev))
        sync(dev)
        ps.start_server(interval_us=30)
        sync(dev)
        g = torch.full((n,), float(rank + 1) * 2.0 ** -10, device=dev)
        x = torch.empty(n, device=dev)
        a = torch.randn(2048, 2048, device=dev)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for t in range(rounds):
            for _ in range(int(skew) * 4):                 # artificial compute skew
                a = torch.tanh(a @ a * 1e-3)
            ps.push(gflat=g)
            ps.pull(x)
        torch.cuda.synchronize()
        mine = time.perf_counter() - t0
        times = [None] * world
        dist.all_gather_object(times, mine)
        sync(dev)
        # let the server threads finish what is still in the slots, then stop them
        deadline = time.time() + 20
        while ps.applied < rounds * world and time.time() < deadline:
            time.sleep(0.005)
        drains = ps.stop_server()
        sync(dev)
        assert ps.applied == rounds * world, (ps.applied, rounds * world)
        assert ps.pushes == rounds
        ps.pull(x)
        torch.cuda.synchronize()
        assert torch.isfinite(x).all()
        if rule == "add":
            want = rounds * sum(r + 1 for r in range(world)) * 2.0 ** -10      # exact in fp32
            assert float(x.min()) == want and float(x.max()) == want, (float(x.min()), float(x.max()), want)
        waits = [None] * world
        dist.all_gather_object(waits, round(ps.max_wait_us, 1))
        if rank == 0:
            print(f"ps_stress {rule}: n={n} rounds={rounds} seconds per rank {[round(t, 4) for t in times]} "
                  f"server drains on rank 0: {drains}, longest flow-control wait per rank (us): {waits}", flush=True)
        ps.check_timeout()
        if world >= 4:
            assert min(times) < 0.7 * max(times), ("the fastest rank was held back by the slowest", times)
        sync(dev)


def case_sync_easgd(rank, world, dev, heap):
    """Synchronous EASGD/EAMSGD as a training path: local Nesterov steps, and every su-th step ONE kernel that
    all-reduces the workers' parameters, applies x -= alpha (x - mean), the look-ahead and the operand refresh.
    Oracle: every rank replays all ranks' trajectories on the CPU (gradients are a fixed function of rank and step)."""
    from mpit_b200.optim.device import DeviceSyncEASGD
    for n, mom in ((10007, 0.9), (351946, 0.0)):          # one-shot and two-shot sizes, EAMSGD and EASGD
        steps, su, lr, alpha = 7, 3, 0.05, 0.4
        grad = lambda r, t: torch.randn(n, generator=torch.Generator().manual_seed(7000 + 100 * r + t)) * 0.1

        class Eng:
            def __init__(self):
                self.wpk = torch.zeros(n, device=dev)
                self.gpk = torch.zeros(n, device=dev)
                self.gmap = None
                self.wmap = None
                self.t = 0

            def load_flat(self, x):
                self.wpk.copy_(x)

            def fwd_bwd(self):
                self.gpk.copy_(grad(rank, self.t))
                self.t += 1

        theta0 = [torch.linspace(-1, 1, n) * (1 + 0.1 * r) for r in range(world)]
        eng = Eng()
        opt = DeviceSyncEASGD(eng, theta0[rank], heap, lr=lr, mva=alpha, su=su, mom=mom, use_graphs=False)
        sync(dev)
        f32 = lambda v: torch.tensor(v, dtype=torch.float32)
        xs = [t.clone() for t in theta0]
        vs = [torch.zeros(n) for _ in range(world)]
        for t in range(steps):
            opt.step()
            # boundary of step t: apply(t-1) happened at the start of this step; then sync or look-ahead
            for r in range(world):
                if t > 0:
                    d = grad(r, t - 1) * f32(-lr)
                    xs[r] = xs[r] + d
                    if mom > 0:
                        vs[r] = vs[r] + d
            if t % su == 0:
                mean = sum(xs) * f32(1.0 / world)
                xs = [x - (x - mean) * f32(alpha) for x in xs]
            if mom > 0:
                for r in range(world):
                    vs[r] = vs[r] * f32(mom)
                    xs[r] = xs[r] + vs[r]
            torch.cuda.synchronize()
            got = eng.wpk.cpu()              # the operand copy holds the look-ahead point of step t
            assert torch.allclose(got, xs[rank], rtol=1e-5, atol=1e-6), (n, t, float((got - xs[rank]).abs().max()))
        opt.coll.check_timeout()
        sync(dev)


def case_msg_device(rank, world, dev, heap):
    """The message API on CUDA tensors over the device transport (msg/device.py): the ring of test.lua, tag matching
    out of order, ANY_SOURCE / ANY_TAG, probe-then-receive, more messages in flight than ring slots, a rendezvous-
    sized message, zero-byte messages, Cancel of an unmatched send and receive, truncation, aio_send / aio_recv under
    the coroutine scheduler, and Allreduce / Bcast (ports of the host-layer cases of tests/mp_cases.py)."""
    from mpit_b200 import msg as mpiT
    from mpit_b200.msg import device as mdev
    mpiT.Init()
    tr = mdev.enable(heap)
    W = mpiT.COMM_WORLD
    f = lambda *v: torch.tensor(v, dtype=torch.float32, device=dev)
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    # -- ring of one float (test.lua:19-38)
    sbuf, rbuf = f(float(rank)), torch.zeros(1, device=dev)
    st = mpiT.Status()
    if rank == 0:
        mpiT.Send(sbuf, 1, mpiT.FLOAT, nxt, 0, W)
        mpiT.Recv(rbuf, 1, mpiT.FLOAT, prv, 0, W, st)
    else:
        mpiT.Recv(rbuf, 1, mpiT.FLOAT, prv, 0, W, st)
        mpiT.Send(sbuf, 1, mpiT.FLOAT, nxt, 0, W)
    assert float(rbuf) == float(prv) and st.source == prv and st.tag == 0 and st.count_bytes == 4
    sync(dev)
    # -- tags out of order + ANY_TAG + probe (rank 0 -> rank 1)
    if rank == 0:
        for tag in (5, 6, 7):
            mpiT.Send(f(tag * 1.0, tag * 2.0), 2, mpiT.FLOAT, 1, tag, W)
    elif rank == 1:
        got = torch.zeros(2, device=dev)
        mpiT.Probe(0, 7, W, st)
        assert st.tag == 7 and st.count_bytes == 8 and mpiT.Get_count(st, mpiT.FLOAT) == 2
        mpiT.Recv(got, 2, mpiT.FLOAT, 0, 7, W, st)
        assert got.tolist() == [7.0, 14.0]
        mpiT.Recv(got, 2, mpiT.FLOAT, 0, mpiT.ANY_TAG, W, st)          # oldest unconsumed first: tag 5
        assert st.tag == 5 and got.tolist() == [5.0, 10.0]
        assert mpiT.Iprobe(0, 6, W, st) and not mpiT.Iprobe(0, 5, W)
        mpiT.Recv(got, 2, mpiT.FLOAT, 0, 6, W)
        assert got.tolist() == [6.0, 12.0]
    sync(dev)
    # -- ANY_SOURCE: everybody -> rank 0
    if rank == 0:
        seen = set()
        got = torch.zeros(1, device=dev)
        for _ in range(world - 1):
            mpiT.Recv(got, 1, mpiT.FLOAT, mpiT.ANY_SOURCE, 11, W, st)
            assert float(got) == float(st.source) * 3
            seen.add(st.source)
        assert seen == set(range(1, world))
    else:
        mpiT.Send(f(rank * 3.0), 1, mpiT.FLOAT, 0, 11, W)
    sync(dev)
    # -- more eager messages in flight than ring slots (flow control), non-blocking on both sides
    n_msg = 3 * tr.K + 1
    reqs = mpiT.Request(n_msg)
    outs = [torch.zeros(64, device=dev) for _ in range(n_msg)]
    for i in range(n_msg):
        mpiT.Irecv(outs[i], 64, mpiT.FLOAT, prv, 20, W, reqs, i)
    sreq = mpiT.Request(n_msg)
    srcs = [torch.full((64,), float(rank * 1000 + i), device=dev) for i in range(n_msg)]
    for i in range(n_msg):
        mpiT.Isend(srcs[i], 64, mpiT.FLOAT, nxt, 20, W, sreq, i)
    mpiT.Waitall(reqs)
    mpiT.Waitall(sreq)
    for i in range(n_msg):
        assert float(outs[i][0]) == prv * 1000 + i and float(outs[i][63]) == prv * 1000 + i, i    # in order
    sync(dev)
    # -- rendezvous-sized message (pulled by the receiver from the sender's tensor), with an offset into it
    n_big = (2 << 20) + 3
    big = torch.arange(n_big + 5, dtype=torch.float32, device=dev) + rank
    land = torch.zeros(n_big, device=dev)
    r1, r2 = mpiT.Request(), mpiT.Request()
    mpiT.Irecv(land, n_big, mpiT.FLOAT, prv, 30, W, r1)
    mpiT.Isend(big[5:], n_big, mpiT.FLOAT, nxt, 30, W, r2)
    while not (mpiT.Test(r1) and mpiT.Test(r2)):
        pass
    assert torch.equal(land, torch.arange(n_big + 5, dtype=torch.float32, device=dev)[5:] + prv)
    sync(dev)
    # -- zero-byte message (descriptor only) and truncation
    z = torch.zeros(4, device=dev)
    if rank == 0:
        mpiT.Send(z, 0, mpiT.FLOAT, 1, 40, W)
        mpiT.Send(z, 4, mpiT.FLOAT, 1, 41, W)
    elif rank == 1:
        mpiT.Recv(z, 0, mpiT.FLOAT, 0, 40, W, st)
        assert st.count_bytes == 0 and st.tag == 40
        try:
            mpiT.Recv(z, 2, mpiT.FLOAT, 0, 41, W, st)
            raise AssertionError("a truncated receive must raise")
        except mpiT.MPIError:
            pass
    sync(dev)
    # -- Cancel: a send nobody receives is withdrawn; so is an unmatched receive (init.lua:52-59 / :94-101)
    if rank == 0:
        rq = mpiT.Request()
        mpiT.Isend(f(1.0), 1, mpiT.FLOAT, 1, 99, W, rq)
        mpiT.Cancel(rq)
        mpiT.Wait(rq, st)
        assert mpiT.Test_cancelled(st)
    rq = mpiT.Request()
    mpiT.Irecv(z, 1, mpiT.FLOAT, nxt, 98, W, rq)
    assert not mpiT.Test(rq)
    mpiT.Cancel(rq)
    mpiT.Wait(rq, st)
    assert mpiT.Test_cancelled(st)
    sync(dev)
    if rank == 1:
        assert not mpiT.Iprobe(0, 99, W), "a cancelled message must not be visible"
    sync(dev)
    # -- coroutine IO over device buffers (init.lua:41-108): aio_send / aio_recv under co_wait
    q = mpiT.Queue()
    a_out = torch.zeros(1000, device=dev)
    a_in = torch.full((1000,), float(rank + 1), device=dev)
    q.push(mpiT.co_execute(lambda: mpiT.aio_recv(a_out, 1000, mpiT.FLOAT, prv, 50, W), ()))
    q.push(mpiT.co_execute(lambda: mpiT.aio_send(a_in, 1000, mpiT.FLOAT, nxt, 50, W), ()))
    mpiT.co_wait(q)
    assert float(a_out.min()) == prv + 1 and float(a_out.max()) == prv + 1
    sync(dev)
    # -- collectives on CUDA tensors
    x = torch.full((5000,), float(rank + 1), device=dev)
    y = torch.zeros(5000, device=dev)
    mpiT.Allreduce(x, y, 5000, mpiT.FLOAT, mpiT.SUM, W)
    torch.cuda.synchronize()
    assert float(y.min()) == world * (world + 1) / 2 and float(y.max()) == world * (world + 1) / 2
    b = torch.full((300,), float(rank), device=dev)
    mpiT.Bcast(b, 300, mpiT.FLOAT, world - 1, W)
    torch.cuda.synchronize()
    assert float(b.min()) == world - 1 and float(b.max()) == world - 1
    sync(dev)
    stats = tr.stats
    assert stats["eager"] > 0 and stats["rendezvous"] > 0 and (rank > 1 or stats["zero"] + stats["probes"] > 0)
    mdev.disable()
    mpiT.Finalize()


def case_pulled_linear(rank, world, dev, heap):
    """Pull fused with the consuming GEMM: weights stay sharded; every rank's GEMM reads them over NVLink -- forward with
    one, two and four M-tiles per CTA, the data gradient (reduction across owners through their transposed shards), the
    weight gradient and its push, and a few steps of training THROUGH the layer on all ranks at once."""
    from mpit_b200.ops.pulled_linear import PulledLinear
    K, N = 256, 64 * world * 2
    gen = torch.Generator().manual_seed(42)
    W = torch.randn(N, K, generator=gen) * 0.05
    b = torch.randn(N, generator=gen) * 0.1
    rpo = N // world
    for M in (200, 512, 700):
        lin = PulledLinear(heap, K, N, M, act=1)
        lin.load_rows(W[rank * rpo:(rank + 1) * rpo].to(dev))
        sync(dev)
        x = torch.randn(M, K, generator=torch.Generator().manual_seed(rank))
        ref = torch.relu(x @ W.t() + b)
        for mt in (0, 1, 2, 4):
            y = lin.forward(x.to(dev), b.to(dev), mt=mt)
            torch.cuda.synchronize()
            lin.check()
            rel = float((y.cpu() - ref).norm() / ref.norm())
            assert rel < 3e-3, (M, mt, rel)
        # the un-fused formulation (gather with in-kernel peer loads, then a local GEMM) gives the same result
        y3 = lin.forward(x.to(dev), b.to(dev), fused=False)
        torch.cuda.synchronize()
        assert float((y3.cpu() - ref).norm() / ref.norm()) < 3e-3
        # backward: dx = dy W (W's rows come from all owners), dW = dy^T x
        y = lin.forward(x.to(dev), b.to(dev))
        dy = torch.randn(M, N, generator=torch.Generator().manual_seed(100 + rank)) * 0.1
        dx = lin.backward(dy.to(dev))
        gw = lin.weight_grad()
        torch.cuda.synchronize()
        lin.check()
        assert float((dx.cpu() - dy @ W).norm() / (dy @ W).norm()) < 3e-3, M
        for mt, ks in ((1, 1), (2, 2), (4, 4), (1, 4 * world)):      # M-tiles per CTA x split-K slices
            dxs = lin.backward(dy.to(dev), mt=mt, ksplit=ks)
            torch.cuda.synchronize()
            lin.check()
            assert float((dxs.cpu() - dy @ W).norm() / (dy @ W).norm()) < 3e-3, (M, mt, ks)
        assert float((gw.cpu() - dy.t() @ x).norm() / (dy.t() @ x).norm()) < 3e-3, M
        sync(dev)
        # push: every rank adds -0.5 * dW into the owners' rows and transposed copies
        lin.push(-0.5)
        sync(dev)
        tot = torch.zeros(N, K)
        for r in range(world):
            xr = torch.randn(M, K, generator=torch.Generator().manual_seed(r))
            dyr = torch.randn(M, N, generator=torch.Generator().manual_seed(100 + r)) * 0.1
            tot += dyr.t() @ xr
        want = (W - 0.5 * tot)[rank * rpo:(rank + 1) * rpo]
        got = lin.weight_rows().cpu()
        assert float((got - want).norm() / want.norm()) < 3e-3
        assert torch.allclose(lin.weight_rows_t().cpu(), got.t(), atol=1e-5), "both layouts stay equal"
        sync(dev)
    # an owner updates its rows; the next forward of every rank sees the new weights without any pull call
    lin.load_rows(lin.weight_rows() * 2.0)
    sync(dev)
    W2 = torch.cat([r for r in _gather_rows(lin, world, dev)])
    y2 = lin.forward(x.to(dev), None)
    torch.cuda.synchronize()
    ref2 = torch.relu(x @ W2.t())
    assert float((y2.cpu() - ref2).norm() / ref2.norm()) < 3e-3
    sync(dev)
    # training through the layer: every rank fits y = relu(x Wt^T) on its own data; the shared, sharded W converges
    Kt, Nt, Mt = 64, 64 * world, 256
    lin = PulledLinear(heap, Kt, Nt, Mt, act=1)
    Wt = torch.randn(Nt, Kt, generator=torch.Generator().manual_seed(7)) * 0.3
    lin.load_rows(torch.zeros(Nt // world, Kt, device=dev) + 0.01)
    sync(dev)
    losses = []
    for step in range(60):
        xs = torch.randn(Mt, Kt, generator=torch.Generator().manual_seed(1000 * rank + step)).to(dev)
        tgt = torch.relu(xs @ Wt.t().to(dev))
        y = lin.forward(xs, None)
        diff = y - tgt
        losses.append(float((diff * diff).mean()))
        dy = diff * (y > 0).float() * (2.0 / (Mt * world))          # d(mean squared error)/d(pre-activation)
        lin.backward(dy)
        lin.weight_grad()
        sync(dev)                        # everybody has read the weights of this step
        lin.push(-0.2)
        sync(dev)
    lin.check()
    assert losses[-1] < 0.2 * losses[0], (losses[0], losses[-1])
    heap.check_timeout()


def case_pulled_mlp(rank, world, dev, heap):
    """A model trained through the pull-fused GEMMs on all ranks at once (mlaunch --model pulled_mlp)."""
    from mpit_b200.workloads import pulled_mlp
    out = pulled_mlp.train(heap, batch=128, steps=120, lr=0.05, hidden=64 * world * 4, barrier=lambda: sync(dev),
                           quiet=rank != 0)
    assert out["loss_last"] < 0.5 * out["loss_first"], out
    assert out["acc_after"] > max(0.5, out["acc_before"] + 0.3), out


def _gather_rows(lin, world, dev):
    rows = [torch.empty(lin.rows_per_owner, lin.K, device=dev) for _ in range(world)]
    dist.all_gather(rows, lin.weight_rows().contiguous())
    return [r.cpu() for r in rows]


def case_window(rank, world, dev, heap):
    """One-sided ring: every rank Puts into its right neighbour, Accumulates into everybody, fences, checks,
    then Gets from its left neighbour (MPI_Put/Accumulate/Get/Win_fence over NVLink)."""
    from mpit_b200.parallel.window import DeviceWindow
    n = 4096 + 3
    win = DeviceWindow(heap, n)
    win.fence()
    mine = torch.full((n,), float(rank + 1), device=dev)
    win.put(mine, (rank + 1) % world)
    win.fence()
    assert torch.equal(win.local, torch.full((n,), float((rank - 1) % world + 1), device=dev))
    win.fence()
    one = torch.ones(n, device=dev)
    for t in range(world):
        win.accumulate(one, t, scale=float(rank + 1))
    win.fence()
    want = float((rank - 1) % world + 1) + world * (world + 1) / 2
    assert torch.allclose(win.local, torch.full((n,), want, device=dev)), (float(win.local[0]), want)
    got = torch.empty(100, device=dev)
    win.get(got, (rank - 1) % world, target_disp=5, count=100)
    win.unlock((rank - 1) % world)
    left = float((rank - 2) % world + 1) + world * (world + 1) / 2
    assert torch.allclose(got, torch.full((100,), left, device=dev))
    win.fence()
    heap.check_timeout()


CASES = {k[5:]: v for k, v in globals().items() if k.startswith("case_")}

if __name__ == "__main__":
    rank, world, dev = setup()
    mode = os.environ.get("MPIT_SYMM_MODE", "auto")
    heap = SymmetricHeap(dev, mode=mode)
    names = sys.argv[1:] or list(CASES)
    for name in names:
        CASES[name](rank, world, dev, heap)
        sync(dev)
        if rank == 0:
            print(f"case {name}: ok (heap mode {heap.mode}, multicast {heap.multicast_ok})", flush=True)
    dist.barrier()
    dist.destroy_process_group()
