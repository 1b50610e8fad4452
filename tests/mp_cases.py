"""Multi-process test bodies, run as ``python tests/mp_cases.py <case>`` under mpit_b200.launch.run.
Each case asserts on values (the reference's smoke scripts only print: /root/reference/test.lua,
test/testreduceall.lua, test/testireduceall.lua, asyncsgd/ptest.lua)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from mpit_b200 import msg as mpiT
from mpit_b200.aio import Queue, IOState, co_execute, co_wait, co_ping, aio_recv, aio_send, signal_INIT, signal_DONE


def case_symm_rendezvous():
    """The symmetric heap's handle exchange over the host message layer (jobs started with launch.run have no
    torch.distributed): object all-gather and barrier."""
    from mpit_b200.parallel import symm
    rank, size = mpiT.get_rank(), mpiT.get_size()
    assert symm._dist_info() == (rank, size)
    objs = symm._all_gather_obj({"rank": rank, "blob": b"x" * (10 + 7 * rank), "pid": os.getpid()})
    assert [o["rank"] for o in objs] == list(range(size))
    assert all(len(o["blob"]) == 10 + 7 * r for r, o in enumerate(objs))
    symm._barrier()


def case_peer_death():
    """A rank that disappears must not leave the others blocked: a receive posted for it, and a send to it that waits
    for the rendezvous acknowledgement, complete with an error once the connection is gone."""
    rank = mpiT.get_rank()
    if rank == 1:
        time.sleep(0.3)
        os._exit(0)                       # no Finalize: the sockets just close
    if rank == 0:
        buf = torch.zeros(4)
        t0 = time.time()
        try:
            mpiT.Recv(buf, 4, mpiT.FLOAT, 1, 7, mpiT.COMM_WORLD)
            raise AssertionError("a receive from a dead rank must fail")
        except mpiT.MPIError:
            pass
        assert time.time() - t0 < 20
        req = mpiT.Request()
        big = torch.zeros(1 << 20)
        try:
            mpiT.Isend(big, big.numel(), mpiT.FLOAT, 1, 8, mpiT.COMM_WORLD, req)
            mpiT.Wait(req)
            raise AssertionError("a send to a dead rank must fail")
        except mpiT.MPIError:
            pass
        print("case peer_death: ok", flush=True)
    os._exit(0)                           # the job cannot run a barrier any more


def case_ring():
    """/root/reference/test.lua:19-38 with an assertion."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    assert size > 1
    s = torch.full((1,), float(rank))
    r = torch.full((1,), float(rank))
    src, dest = (rank - 1) % size, (rank + 1) % size
    st = mpiT.Status()
    if rank == 0:
        mpiT.Send(s, 1, mpiT.FLOAT, dest, 0, mpiT.COMM_WORLD)
        mpiT.Recv(r, 1, mpiT.FLOAT, src, 0, mpiT.COMM_WORLD, st)
    else:
        mpiT.Recv(r, 1, mpiT.FLOAT, src, 0, mpiT.COMM_WORLD, st)
        mpiT.Send(s, 1, mpiT.FLOAT, dest, 0, mpiT.COMM_WORLD)
    assert float(r[0]) == float(src), (rank, r)
    assert st.source == src and st.tag == 0 and mpiT.Get_count(st, mpiT.FLOAT) == 1


def case_allreduce():
    """/root/reference/test/testreduceall.lua:19-31: every rank seeds 123, so the sum is size * a."""
    size = mpiT.get_size()
    megs = float(os.environ.get("MEGS", "1"))
    n = int(megs * 1024 * 1024 / 4)
    torch.manual_seed(123)
    a = torch.rand(n)
    out = torch.zeros(n)
    t0 = mpiT.Wtime()
    mpiT.Allreduce(a, out, n, mpiT.FLOAT, mpiT.SUM, mpiT.COMM_WORLD)
    dt = mpiT.Wtime() - t0
    assert torch.allclose(out, a * size, rtol=1e-6), float((out - a * size).abs().max())
    assert dt >= 0 and mpiT.Wtick() > 0
    # other ops / dtypes
    x = torch.tensor([mpiT.get_rank() + 1], dtype=torch.int64)
    y = torch.zeros(1, dtype=torch.int64)
    mpiT.Allreduce(x, y, 1, mpiT.LONG, mpiT.MAX)
    assert int(y) == size
    mpiT.Allreduce(x, y, 1, mpiT.LONG, mpiT.PROD)
    import math
    assert int(y) == math.factorial(size)


def case_iallreduce():
    """/root/reference/test/testireduceall.lua:27-39: Iallreduce -> Test -> Wait -> Test."""
    size = mpiT.get_size()
    n = 1 << 18
    torch.manual_seed(123)
    a = torch.rand(n)
    out = torch.zeros(n)
    req, st = mpiT.Request(), mpiT.Status()
    mpiT.Iallreduce(a, out, n, mpiT.FLOAT, mpiT.SUM, mpiT.COMM_WORLD, req)
    first = mpiT.Test(req, st)          # may or may not be complete yet
    if not first:
        assert not req.is_null()
        mpiT.Wait(req, st)
    assert req.is_null()
    assert mpiT.Test(req, st) is True   # a null request tests complete
    assert torch.allclose(out, a * size, rtol=1e-6)
    # two overlapping non-blocking all-reduces on one communicator do not cross
    b, c = torch.full((8,), 1.0), torch.full((8,), 2.0)
    ob, oc = torch.zeros(8), torch.zeros(8)
    r2 = mpiT.Request(2)
    mpiT.Iallreduce(b, ob, 8, mpiT.FLOAT, mpiT.SUM, mpiT.COMM_WORLD, r2, 0)
    mpiT.Iallreduce(c, oc, 8, mpiT.FLOAT, mpiT.SUM, mpiT.COMM_WORLD, r2, 1)
    mpiT.Waitall(r2)
    assert float(ob[0]) == size and float(oc[0]) == 2 * size


def case_matching():
    """Tag matching, wildcards, probe, zero-byte messages, truncation, non-overtaking order."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    W = mpiT.COMM_WORLD
    if rank == 0:
        for tag in (7, 5, 9):
            mpiT.Send(torch.full((4,), float(tag)), 4, mpiT.FLOAT, 1, tag, W)
        mpiT.Send(None, 0, mpiT.FLOAT, 1, 11, W)                      # zero-byte header message
        for k in range(5):                                            # same tag: must stay ordered
            mpiT.Send(torch.full((1,), float(k)), 1, mpiT.FLOAT, 1, 3, W)
        mpiT.Send(torch.arange(8, dtype=torch.float32), 8, mpiT.FLOAT, 1, 13, W)   # will be truncated
    elif rank == 1:
        st = mpiT.Status()
        buf = torch.zeros(4)
        for tag in (5, 9, 7):                                         # receive out of send order
            mpiT.Recv(buf, 4, mpiT.FLOAT, 0, tag, W, st)
            assert float(buf[0]) == tag and st.tag == tag
        mpiT.Probe(mpiT.ANY_SOURCE, 11, W, st)
        assert st.source == 0 and st.count_bytes == 0
        mpiT.Recv(None, 0, mpiT.FLOAT, 0, 11, W, st)
        one = torch.zeros(1)
        for k in range(5):
            mpiT.Recv(one, 1, mpiT.FLOAT, mpiT.ANY_SOURCE, mpiT.ANY_TAG, W, st)
            assert float(one) == k and st.tag == 3
        while not mpiT.Iprobe(0, 13, W, st):
            time.sleep(0.001)
        assert mpiT.Get_count(st, mpiT.FLOAT) == 8
        small = torch.zeros(4)
        try:
            mpiT.Recv(small, 4, mpiT.FLOAT, 0, 13, W, st)
            raise AssertionError("truncation not reported")
        except mpiT.MPIError as e:
            assert e.code == mpiT.ERR_TRUNCATE
        assert torch.equal(small, torch.arange(4, dtype=torch.float32))
        assert not mpiT.Iprobe(mpiT.ANY_SOURCE, mpiT.ANY_TAG, W)
    mpiT.Barrier()
    s = mpiT.stats()
    assert s["unexpected_now"] == 0 and s["posted_now"] == 0, s
    # type checking the reference does not have (lua-mpi.h:70-76)
    try:
        mpiT.Send(torch.zeros(2, dtype=torch.float64), 2, mpiT.FLOAT, 0, 1, W)
        raise AssertionError("dtype mismatch accepted")
    except TypeError:
        pass
    try:
        mpiT.Send(torch.zeros(2), 3, mpiT.FLOAT, 0, 1, W)
        raise AssertionError("overlong count accepted")
    except ValueError:
        pass


def case_cancel():
    """Cancel has real semantics (the reference leaks a receive, /root/reference/init.lua:88-101)."""
    rank = mpiT.get_rank()
    W = mpiT.COMM_WORLD
    req, st = mpiT.Request(), mpiT.Status()
    buf = torch.zeros(16)
    mpiT.Irecv(buf, 16, mpiT.FLOAT, mpiT.ANY_SOURCE, 99, W, req)     # nobody sends tag 99
    assert not mpiT.Test(req, st)
    mpiT.Cancel(req)
    mpiT.Wait(req, st)
    assert mpiT.Test_cancelled(st) and req.is_null()
    assert mpiT.stats()["posted_now"] == 0
    # aio_recv / aio_send under a dropped io flag
    state = IOState()
    log = []

    def body():
        yield signal_INIT
        got = yield from aio_recv(buf, 16, mpiT.FLOAT, mpiT.ANY_SOURCE, 98, W, state)
        log.append(got)
        yield signal_DONE

    q = Queue()
    q.push(co_execute(body))
    for _ in range(5):
        assert co_ping(q)
    state["io"] = False
    co_wait(q)
    assert log == [False] and mpiT.stats()["posted_now"] == 0
    mpiT.Barrier()


def case_comm():
    rank, size = mpiT.get_rank(), mpiT.get_size()
    sub = mpiT.Comm_split(mpiT.COMM_WORLD, rank % 2, -rank)           # reverse order inside each half
    members = [r for r in range(size) if r % 2 == rank % 2]
    assert mpiT.Comm_size(sub) == len(members)
    assert mpiT.Comm_rank(sub) == sorted(members, reverse=True).index(rank)
    x, y = torch.tensor([float(rank)]), torch.zeros(1)
    mpiT.Allreduce(x, y, 1, mpiT.FLOAT, mpiT.SUM, sub)
    assert float(y) == sum(members)
    dup = mpiT.Comm_dup(mpiT.COMM_WORLD)
    assert mpiT.Comm_size(dup) == size and mpiT.Comm_rank(dup) == rank
    # same tag on two communicators does not cross-match
    if rank == 0:
        mpiT.Send(torch.tensor([1.0]), 1, mpiT.FLOAT, 1, 4, dup)
        mpiT.Send(torch.tensor([2.0]), 1, mpiT.FLOAT, 1, 4, mpiT.COMM_WORLD)
    if rank == 1:
        a = torch.zeros(1)
        mpiT.Recv(a, 1, mpiT.FLOAT, 0, 4, mpiT.COMM_WORLD)
        assert float(a) == 2.0
        mpiT.Recv(a, 1, mpiT.FLOAT, 0, 4, dup)
        assert float(a) == 1.0
    mpiT.Comm_free(sub)
    mpiT.Comm_free(dup)
    assert mpiT.Comm_size(mpiT.COMM_SELF) == 1


def case_collectives():
    rank, size = mpiT.get_rank(), mpiT.get_size()
    b = torch.arange(5, dtype=torch.float32) if rank == 1 else torch.zeros(5)
    mpiT.Bcast(b, 5, mpiT.FLOAT, 1)
    assert torch.equal(b, torch.arange(5, dtype=torch.float32))
    mine = torch.tensor([rank, rank * 10], dtype=torch.int32)
    allg = torch.zeros(2 * size, dtype=torch.int32)
    mpiT.Allgather(mine, 2, mpiT.INT, allg)
    assert allg.tolist() == [v for r in range(size) for v in (r, r * 10)]
    g = torch.zeros(2 * size, dtype=torch.int32)
    mpiT.Gather(mine, 2, mpiT.INT, g if rank == 0 else None, 0)
    if rank == 0:
        assert torch.equal(g, allg)
    sc = torch.zeros(2, dtype=torch.int32)
    mpiT.Scatter(allg if rank == 0 else None, 2, mpiT.INT, sc, 0)
    assert torch.equal(sc, mine)
    a2a_in = torch.tensor([rank * 100 + d for d in range(size)], dtype=torch.int64)
    a2a_out = torch.zeros(size, dtype=torch.int64)
    mpiT.Alltoall(a2a_in, 1, mpiT.LONG, a2a_out)
    assert a2a_out.tolist() == [s * 100 + rank for s in range(size)]
    red = torch.zeros(3, dtype=torch.float64)
    mpiT.Reduce(torch.full((3,), float(rank), dtype=torch.float64), red, 3, mpiT.DOUBLE, mpiT.SUM, size - 1)
    if rank == size - 1:
        assert float(red[0]) == sum(range(size))
    scan = torch.zeros(1)
    mpiT.Scan(torch.tensor([1.0 + rank]), scan, 1, mpiT.FLOAT, mpiT.SUM)
    assert float(scan) == sum(1.0 + r for r in range(rank + 1))
    rs = torch.zeros(2)
    mpiT.Reduce_scatter_block(torch.ones(2 * size) * (rank + 1), rs, 2, mpiT.FLOAT, mpiT.SUM)
    assert float(rs[0]) == sum(r + 1 for r in range(size))
    sb, rb = torch.tensor([float(rank)]), torch.zeros(1)
    mpiT.Sendrecv(sb, 1, mpiT.FLOAT, (rank + 1) % size, 2, rb, 1, mpiT.FLOAT, (rank - 1) % size, 2)
    assert float(rb) == (rank - 1) % size
    assert len(mpiT.Get_processor_name()) > 0
    obj = {"a": 1.5, "t": torch.arange(6, dtype=torch.float32).view(2, 3), "s": "hello"}
    back = mpiT.deserialize(mpiT.serialize(obj))
    assert back["a"] == 1.5 and back["s"] == "hello" and torch.equal(back["t"], obj["t"])


CASES = {k[5:]: v for k, v in globals().items() if k.startswith("case_")}

if __name__ == "__main__":
    from tests import mp_cases_ps, mp_cases_extras, mp_cases_chapters
    CASES.update(mp_cases_ps.CASES)
    CASES.update(mp_cases_extras.CASES)
    CASES.update(mp_cases_chapters.CASES)
    name = sys.argv[1]
    mpiT.Init()
    try:
        CASES[name]()
        mpiT.Barrier()
    finally:
        mpiT.Finalize()
    if mpiT.__dict__["_state"]["rank"] == 0:
        print(f"case {name}: ok")
