"""Parameter-server multi-process cases (registered into tests.mp_cases.CASES)."""
import math
import os

import torch

from mpit_b200 import msg as mpiT
from mpit_b200 import ps
from mpit_b200.optim import reference as R
from mpit_b200.parallel.center import partition


def _roles(layout="halves", **extra):
    rank, size = mpiT.get_rank(), mpiT.get_size()
    return rank, size, ps.make_conf(rank, size, layout, **extra)


def case_ps_protocol():
    """Partition with remainder, first-client seeding, per-client gradient slots, conservation of
    pushed updates, pull, stop votes (asyncsgd/pserver.lua, pclient.lua; asyncsgd/ptest.lua shape)."""
    rank, size, conf = _roles("halves")
    N = 1003
    ns, nc = len(conf["sranks"]), len(conf["cranks"])
    if rank in conf["sranks"]:
        srv = ps.pServer(conf)
        srv.start()
        off, sz = partition(N, ns)[conf["sranks"].index(rank)]
        assert (srv.offset, srv.size) == (off, sz)
        assert set(srv.g.keys()) == set(conf["cranks"])
        # centre = seed (first client's theta) + sum over clients and rounds of pushed grads
        ci = torch.arange(N, dtype=torch.float32)[off:off + sz]
        seed = ci * 0.5 + conf["cranks"][0]
        pushed = sum(3 * (c + 1) * 0.25 for c in conf["cranks"])
        assert torch.allclose(srv.p, seed + pushed), float((srv.p - seed - pushed).abs().max())
        assert srv.n_updates == 3 * nc and srv.state["iostop"] == nc
    else:
        theta = torch.arange(N, dtype=torch.float32) * 0.5 + rank
        grad = torch.full((N,), (rank + 1) * 0.25)
        pc = ps.pClient(conf)
        pc.start(theta, grad)
        for _ in range(3):
            pc.async_send_grad()
            pc.async_recv_param()
            pc.wait()
        # after the last round theta is some consistent-per-shard snapshot of the centre:
        seed = torch.arange(N, dtype=torch.float32) * 0.5 + conf["cranks"][0]
        d = theta - seed
        assert float(d.min()) >= 0.25 * 3 - 1e-5 or nc > 1   # own pushes are in (single client)
        assert float(d.max()) <= sum(3 * (c + 1) * 0.25 for c in conf["cranks"]) + 1e-4
        pc.stop()
        assert mpiT.stats()["posted_now"] == 0


def case_ps_eamsgd():
    """EAMSGD over the host PS with a quadratic objective: workers and centre converge to the optimum."""
    rank, size, conf = _roles("parity")
    N = 257
    target = torch.linspace(-1, 1, N)
    if rank in conf["sranks"]:
        srv = ps.pServer(conf)
        srv.start()
        off, sz = partition(N, len(conf["sranks"]))[conf["sranks"].index(rank)]
        assert float((srv.p - target[off:off + sz]).abs().max()) < 0.05
    else:
        torch.manual_seed(rank)
        w = torch.randn(N)
        grad = torch.zeros(N)
        pc = ps.pClient(conf)
        pc.start(w, grad)
        p = len(conf["cranks"])
        cfg = {"lr": 0.05, "mom": 0.9, "mva": 0.9 / p, "su": 2, "pclient": pc}

        def opfunc(x):
            return float(0.5 * ((x - target) ** 2).sum()), (x - target).clone()

        for _ in range(300):
            R.eamsgd(opfunc, w, cfg)
        pc.stop()
        assert float((w - target).abs().max()) < 0.05
        assert cfg["dusync"] >= 0.0


def case_ps_downpour():
    rank, size, conf = _roles("parity")
    N = 100
    target = torch.linspace(0, 1, N)
    if rank in conf["sranks"]:
        srv = ps.pServer(conf)
        srv.start()
        off, sz = partition(N, len(conf["sranks"]))[conf["sranks"].index(rank)]
        assert float((srv.p - target[off:off + sz]).abs().max()) < 0.05
    else:
        w = torch.zeros(N)
        grad = torch.zeros(N)
        pc = ps.pClient(conf)
        pc.start(w, grad)
        cfg = {"lr": 0.05, "su": int(os.environ.get("SU", "1")), "pclient": pc}

        def opfunc(x):
            g = (x - target).clone()
            if cfg["su"] == 1:
                grad.copy_(g)        # su==1 pushes the client's grad storage
                return 0.0, grad
            return 0.0, g

        for _ in range(400):
            R.downpour(opfunc, w, cfg)
        pc.stop()
        assert float((w - target).abs().max()) < 0.08


def case_ps_server_adam():
    """Server-side Adam (BiCNN/pserver.lua:140-155) against the single-process oracle: one server, one
    client, deterministic gradient sequence."""
    rank, size, conf = _roles("halves", opt={"optimization": "adam", "modeAdam": "global", "lrAdam": 0.01,
                                             "beta1Adam": 0.9, "beta2Adam": 0.999, "epsilonAdam": 1e-8,
                                             "stepDivAdam": 2})
    N = 64
    gen = torch.Generator().manual_seed(0)
    grads = [torch.randn(N, generator=gen) for _ in range(5)]
    w0 = torch.randn(N, generator=gen)
    if rank in conf["sranks"]:
        srv = ps.pServer(conf)
        srv.start()
        p = w0.clone()
        st = R.rule_state("adam", p)
        hp = dict(lr=0.01, beta1=0.9, beta2=0.999, epsilon=1e-8, step_div=2)
        for g in grads:
            R.apply_rule("adam", p, g, st, hp)
        assert torch.equal(srv.p, p)
    else:
        w = w0.clone()
        grad = torch.zeros(N)
        pc = ps.pClient(conf)
        pc.start(w, grad)
        it = iter(grads)
        cfg = {"pclient": pc, "su": 1, "mode": "global"}

        def opfunc(x):
            grad.copy_(next(it))
            return 0.0, grad

        for _ in grads:
            R.adam(opfunc, w, cfg)
        # Each step queues its push and its pull together, and the server serves them from separate coroutines with no
        # lock between them ("expect inconsistent read", pserver.lua:81): the pull of step t may or may not contain
        # gradient t.  Every push has been acknowledged by now, so one more pull is the server's final state.
        pc.async_recv_param()
        pc.wait()
        pc.stop()
        p = w0.clone()
        st = R.rule_state("adam", p)
        hp = dict(lr=0.01, beta1=0.9, beta2=0.999, epsilon=1e-8, step_div=2)
        for g in grads:
            R.apply_rule("adam", p, g, st, hp)
        assert torch.equal(w, p)    # the pulled parameters are the server's


def case_ps_single_mode():
    """Single-worker + shadow reader (BiCNN/pserver.lua:280-289, optim-*-single.lua, bicnn.lua:580-596)."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    assert size == 3
    conf = {"rank": rank, "world": mpiT.COMM_WORLD, "sranks": [0], "cranks": [1, 2], "size": size,
            "opt": {"singlemode": True, "single_writer": 1, "single_readers": [2]}}
    N = 50
    if rank == 0:
        ps.pServer(conf).start()
    elif rank == 1:
        w = torch.ones(N)
        pc = ps.pClient(conf)
        pc.start(w, torch.zeros(N))
        cfg = {"pclient": pc, "lr": 0.1, "epsilon": 1e-10}
        for _ in range(3):
            R.adagradsingle(lambda x: (0.0, torch.ones(N)), w, cfg)
        mpiT.Send(w, N, mpiT.FLOAT, 2, 77)
        pc.stop()
    else:
        mirror = torch.zeros(N)
        pc = ps.pClient(conf)
        pc.start(mirror, torch.zeros(N))
        final = torch.zeros(N)
        mpiT.Recv(final, N, mpiT.FLOAT, 1, 77)    # the writer has pushed everything by now
        pc.async_recv_param()
        pc.wait()
        assert torch.equal(mirror, final)
        pc.stop()


def case_ptest():
    """asyncsgd/ptest.lua: bidirectional PS bandwidth (size reduced), prints MB/s with a real timer."""
    rank, size, conf = _roles("halves")
    ssize = int(os.environ.get("SSIZE", str(1 << 18)))
    T = int(os.environ.get("ROUNDS", "5"))
    if rank in conf["sranks"]:
        ps.pServer(conf).start()
    else:
        theta, grad = torch.zeros(ssize), torch.zeros(ssize)
        pc = ps.pClient(conf)
        pc.start(theta, grad)
        t0 = mpiT.Wtime()
        for _ in range(T):
            pc.async_recv_param()
            pc.async_send_grad()
            pc.wait()
        dt = mpiT.Wtime() - t0
        bw = 2 * T * ssize * 4 / dt / 1024 / 1024
        assert math.isfinite(bw) and bw > 0
        print(f"rank {rank} bandwidth(bi-direction) is {bw:.1f} MBytes/sec")
        pc.stop()


CASES = {k[5:]: v for k, v in globals().items() if k.startswith("case_")}
