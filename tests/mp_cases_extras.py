"""Multi-process bodies for mpit_b200.msg.extras (the wider MPI surface of /root/reference/mpifuncs.c)."""
import torch

from mpit_b200 import msg as mpiT


def case_x_vector_collectives():
    """Gatherv / Scatterv / Allgatherv / Alltoallv / Exscan / Reduce_scatter with uneven counts."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    counts = [r + 1 for r in range(size)]
    displs = [sum(counts[:r]) for r in range(size)]
    total = sum(counts)
    mine = torch.full((counts[rank],), float(rank))
    want = torch.cat([torch.full((c,), float(r)) for r, c in enumerate(counts)])
    out = torch.zeros(total)
    mpiT.Gatherv(mine, counts[rank], mpiT.FLOAT, out, counts, displs, 1)
    if rank == 1:
        assert torch.equal(out, want)
    out.zero_()
    mpiT.Allgatherv(mine, counts[rank], mpiT.FLOAT, out, counts, displs)
    assert torch.equal(out, want)
    back = torch.zeros(counts[rank])
    mpiT.Scatterv(want if rank == 0 else None, counts, displs, mpiT.FLOAT, back, counts[rank], 0)
    assert torch.equal(back, mine)
    # alltoallv: rank r sends (p+1) copies of r*10+p to rank p
    sc = [p + 1 for p in range(size)]
    sd = [sum(sc[:p]) for p in range(size)]
    sbuf = torch.cat([torch.full((p + 1,), float(rank * 10 + p)) for p in range(size)])
    rc = [rank + 1] * size
    rd = [p * (rank + 1) for p in range(size)]
    rbuf = torch.zeros(sum(rc))
    mpiT.Alltoallv(sbuf, sc, sd, mpiT.FLOAT, rbuf, rc, rd)
    assert torch.equal(rbuf, torch.cat([torch.full((rank + 1,), float(p * 10 + rank)) for p in range(size)]))
    # exscan
    x = torch.tensor([float(rank + 1)])
    y = torch.full((1,), -7.0)
    mpiT.Exscan(x, y, 1, mpiT.FLOAT, mpiT.SUM)
    assert float(y) == (-7.0 if rank == 0 else rank * (rank + 1) / 2)
    # reduce_scatter with uneven counts
    full = torch.arange(total, dtype=torch.float32) + rank
    piece = torch.zeros(counts[rank])
    mpiT.Reduce_scatter(full, piece, counts, mpiT.FLOAT, mpiT.SUM)
    base = torch.arange(total, dtype=torch.float32) * size + size * (size - 1) / 2
    assert torch.equal(piece, base[displs[rank]:displs[rank] + counts[rank]])


def case_x_requests():
    """Persistent requests, Testany / Waitsome, send modes, synchronous send handshake."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    right, left = (rank + 1) % size, (rank - 1) % size
    s, r = torch.zeros(4), torch.zeros(4)
    req = mpiT.Request(2)
    mpiT.Recv_init(r, 4, mpiT.FLOAT, left, 7, mpiT.COMM_WORLD, req, 0)
    mpiT.Send_init(s, 4, mpiT.FLOAT, right, 7, mpiT.COMM_WORLD, req, 1)
    for it in range(3):                                  # the same request pair restarted three times
        s.fill_(rank * 100 + it)
        mpiT.Startall(req)
        done = set()
        while len(done) < 2:
            done.update(mpiT.Waitsome(req))
        assert torch.equal(r, torch.full((4,), float(left * 100 + it)))
    assert mpiT.Testany(req) == (mpiT.UNDEFINED, True)   # nothing active any more
    mpiT.Request_free(req, 0)
    mpiT.Request_free(req, 1)
    for send in (mpiT.Bsend, mpiT.Rsend):
        if rank == 0:
            send(torch.tensor([3.5]), 1, mpiT.FLOAT, 1, 9)
        elif rank == 1:
            b = torch.zeros(1)
            mpiT.Recv(b, 1, mpiT.FLOAT, 0, 9)
            assert float(b) == 3.5
    if rank == 0:
        mpiT.Ssend(torch.tensor([1.25]), 1, mpiT.FLOAT, 1, 11)
    elif rank == 1:
        b = torch.zeros(1)
        mpiT.Recv(b, 1, mpiT.FLOAT, 0, 11)
        mpiT.Ssend_ack(0)
        assert float(b) == 1.25
    q = mpiT.Request(1)
    mpiT.Issend(torch.tensor([2.0]), 1, mpiT.FLOAT, right, 12, mpiT.COMM_WORLD, q)
    b = torch.zeros(1)
    mpiT.Recv(b, 1, mpiT.FLOAT, left, 12)
    mpiT.Wait(q)
    assert float(b) == 2.0


def case_x_groups_ops():
    """Groups, Comm_create / Comm_compare, user-defined and XOR reduction ops, error strings."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    g = mpiT.Comm_group(mpiT.COMM_WORLD)
    assert mpiT.Group_size(g) == size and mpiT.Group_rank(g) == rank
    evens = mpiT.Group_incl(g, list(range(0, size, 2)))
    odds = mpiT.Group_excl(g, list(range(0, size, 2)))
    assert mpiT.Group_compare(mpiT.Group_union(evens, odds), g) in (mpiT.IDENT, mpiT.SIMILAR)
    assert mpiT.Group_size(mpiT.Group_intersection(evens, odds)) == 0
    assert mpiT.Group_translate_ranks(evens, [0], g) == [0]
    sub = mpiT.Comm_create(mpiT.COMM_WORLD, evens)
    if rank % 2 == 0:
        assert mpiT.Comm_size(sub) == (size + 1) // 2 and mpiT.Comm_rank(sub) == rank // 2
        assert mpiT.Comm_compare(sub, sub) == mpiT.IDENT
        mpiT.Comm_set_name(sub, "evens")
        assert mpiT.Comm_get_name(sub) == "evens"
    else:
        assert sub == mpiT.COMM_NULL
    dup = mpiT.Comm_dup(mpiT.COMM_WORLD)
    assert mpiT.Comm_compare(dup, mpiT.COMM_WORLD) == mpiT.CONGRUENT
    # user op: elementwise "keep the value with the larger magnitude"
    def absmax(invec, inout):
        take = invec.abs() > inout.abs()
        inout[take] = invec[take]
    op = mpiT.Op_create(absmax, True)
    x = torch.tensor([float(rank) - 1.5, -float(rank)])
    y = torch.zeros(2)
    mpiT.Allreduce(x, y, 2, mpiT.FLOAT, op)
    assert y.tolist() == [float(size - 1) - 1.5, -float(size - 1)], y
    z = torch.zeros(2)
    mpiT.Reduce(x, z, 2, mpiT.FLOAT, op, 1)
    if rank == 1:
        assert torch.equal(z, y)
    mpiT.Op_free(op)
    a = torch.tensor([1 << rank], dtype=torch.int64)
    b = torch.zeros(1, dtype=torch.int64)
    mpiT.Allreduce(a, b, 1, mpiT.LONG, mpiT.BXOR)
    assert int(b) == (1 << size) - 1
    assert "TRUNCATE" in mpiT.Error_string(mpiT.ERR_TRUNCATE) and mpiT.Get_version()[0] >= 3
    assert mpiT.Type_size(mpiT.DOUBLE) == 8 and mpiT.Query_thread() == mpiT.THREAD_MULTIPLE


def case_x_window():
    """MPI_Win_create / Put / Accumulate / Get on host memory: fence epochs, then passive-target lock / unlock."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    n = 16
    base = torch.zeros(n)
    win = mpiT.Win_create(base, n * 4, 4, None, mpiT.COMM_WORLD)
    mpiT.Win_fence(0, win)
    right, left = (rank + 1) % size, (rank - 1) % size
    mpiT.Put(torch.full((4,), float(rank + 1)), 4, mpiT.FLOAT, right, 2, win)
    for t in range(size):
        mpiT.Accumulate(torch.ones(3), 3, mpiT.FLOAT, t, 8, mpiT.SUM, win)
    mpiT.Win_fence(0, win)
    want = torch.zeros(n)
    want[2:6] = float(left + 1)
    want[8:11] = float(size)
    assert torch.equal(base, want), (rank, base)
    got = torch.zeros(4)
    mpiT.Get(got, 4, mpiT.FLOAT, left, 2, win)
    mpiT.Win_fence(0, win)
    assert torch.equal(got, torch.full((4,), float((left - 1) % size + 1)))
    # passive target: rank 0 takes no part.  Every rank does a read-modify-write of element 0 of rank 0's window,
    # which is only correct if the exclusive lock really excludes the others; then everybody reads under a shared lock
    base.zero_()
    mpiT.Barrier()
    if rank != 0:
        mpiT.Win_lock(mpiT.LOCK_EXCLUSIVE, 0, 0, win)
        cur = torch.zeros(1)
        mpiT.Get(cur, 1, mpiT.FLOAT, 0, 0, win)
        mpiT.Put(cur + rank + 1, 1, mpiT.FLOAT, 0, 0, win)
        mpiT.Accumulate(torch.ones(2), 2, mpiT.FLOAT, 0, 4, mpiT.SUM, win)
        mpiT.Win_unlock(0, win)
        try:
            mpiT.Win_unlock(0, win)
            raise AssertionError("unlock without lock")
        except mpiT.MPIError as e:
            assert e.code == mpiT.ERR_RMA_SYNC
    mpiT.Barrier()                       # every unlock has returned, i.e. every update is applied at rank 0
    mpiT.Win_lock(mpiT.LOCK_SHARED, 0, 0, win)
    seen = torch.zeros(6)
    mpiT.Get(seen, 6, mpiT.FLOAT, 0, 0, win)
    mpiT.Win_unlock(0, win)
    assert float(seen[0]) == sum(r + 1 for r in range(1, size)) and seen[4:6].tolist() == [size - 1.0] * 2, seen
    # locks on two targets at once, operations to each land where they should
    mpiT.Barrier()
    mpiT.Win_lock(mpiT.LOCK_SHARED, left, 0, win)
    mpiT.Win_lock(mpiT.LOCK_SHARED, right, 0, win)
    mpiT.Accumulate(torch.tensor([100.0]), 1, mpiT.FLOAT, left, 12, mpiT.SUM, win)
    mpiT.Accumulate(torch.tensor([1.0]), 1, mpiT.FLOAT, right, 12, mpiT.SUM, win)
    mpiT.Win_unlock(right, win)
    mpiT.Win_unlock(left, win)
    mpiT.Barrier()
    assert float(base[12]) == 101.0, base
    mpiT.Win_free(win)


CASES = {k[5:]: v for k, v in globals().items() if k.startswith("case_")}
