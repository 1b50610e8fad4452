"""Multi-process bodies for the MPI chapters in mpit_b200/msg/{datatypes,fileio,topo,attrs,rma,dynproc}.py."""
import os
import socket
import tempfile

import numpy as np
import torch

from mpit_b200 import msg as mpiT


def _shared_tmp(name):
    """A path every rank of this launch agrees on."""
    job = os.environ.get("MPIT_PORT", os.environ.get("MASTER_PORT", "0"))
    return os.path.join(tempfile.gettempdir(), f"mpit-test-{os.getuid()}-{job}-{name}")


def case_c_derived_messages():
    """Send / Recv / Isend / Irecv / Bcast / Sendrecv_replace / Alltoallw with committed derived datatypes."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    right, left = (rank + 1) % size, (rank - 1) % size
    col = mpiT.Type_commit(mpiT.Type_create_resized(mpiT.Type_vector(4, 1, 5, mpiT.FLOAT), 0, 4))
    a = torch.arange(20, dtype=torch.float32).view(4, 5) + 100 * rank
    got = torch.zeros(4, 5)
    # my column 1 and 2 go to the right neighbour's columns 3 and 4 (strided on both sides)
    req = mpiT.Request(2)
    mpiT.Irecv(got.view(-1)[3:], 2, col, left, 5, mpiT.COMM_WORLD, req, 0)
    mpiT.Isend(a.view(-1)[1:], 2, col, right, 5, mpiT.COMM_WORLD, req, 1)
    st = mpiT.Status()
    mpiT.Wait(req, st, 0)
    mpiT.Wait(req, None, 1)
    assert mpiT.Get_count(st, col) == 2 and mpiT.Get_elements(st, col) == 8
    want = torch.zeros(4, 5)
    want[:, 3:5] = (torch.arange(20, dtype=torch.float32).view(4, 5) + 100 * left)[:, 1:3]
    assert torch.equal(got, want), (rank, got)
    # derived type on one side, contiguous floats on the other
    flat = torch.zeros(4)
    mpiT.Sendrecv(a.view(-1)[2:], 1, col, right, 6, flat, 4, mpiT.FLOAT, left, 6)
    assert torch.equal(flat, torch.arange(20, dtype=torch.float32).view(4, 5)[:, 2] + 100 * left)
    # a short message leaves the rest of a strided receive buffer untouched
    keep = torch.full((4, 5), -1.0)
    if rank == 0:
        mpiT.Send(torch.tensor([7.0, 8.0]), 2, mpiT.FLOAT, 1, 9)
    elif rank == 1:
        mpiT.Recv(keep, 1, col, 0, 9, mpiT.COMM_WORLD, st)
        assert keep[:, 0].tolist() == [7.0, 8.0, -1.0, -1.0] and mpiT.Get_count(st, col) == mpiT.UNDEFINED
        assert mpiT.Get_elements(st, col) == 2
    # broadcast of a sub-block
    blockt = mpiT.Type_commit(mpiT.Type_create_subarray([4, 5], [2, 2], [1, 3], mpiT.ORDER_C, mpiT.FLOAT))
    b = a.clone() if rank == 2 % size else torch.zeros(4, 5)
    mpiT.Bcast(b, 1, blockt, 2 % size)
    ref = torch.arange(20, dtype=torch.float32).view(4, 5) + 100 * (2 % size)
    assert torch.equal(b[1:3, 3:5], ref[1:3, 3:5])
    if rank != 2 % size:
        assert float(b.sum()) == float(ref[1:3, 3:5].sum())
    # Sendrecv_replace around the ring
    tok = torch.tensor([float(rank)])
    mpiT.Sendrecv_replace(tok, 1, mpiT.FLOAT, right, 3, left, 3)
    assert float(tok) == left
    # Alltoallw: to peer p I send a strided pair out of row p; from peer p I receive two contiguous floats at byte p*8
    pair = mpiT.Type_commit(mpiT.Type_vector(2, 1, 2, mpiT.FLOAT))
    src = (torch.arange(size * 4, dtype=torch.float32) + 1000 * rank).view(size, 4)
    dst = torch.zeros(size * 2)
    mpiT.Alltoallw(src, [1] * size, [p * 16 for p in range(size)], [pair] * size,
                   dst, [2] * size, [p * 8 for p in range(size)], [mpiT.FLOAT] * size)
    want = torch.cat([torch.tensor([1000.0 * p + rank * 4, 1000.0 * p + rank * 4 + 2]) for p in range(size)])
    assert torch.equal(dst, want), (rank, dst, want)


def case_c_maxloc_and_wide_types():
    """MAXLOC / MINLOC over the pair types, reductions of the predefined types without a native wire id, and a
    contiguous derived type of floats."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    pair = np.dtype([("v", np.float32), ("i", np.int32)])
    mine = np.zeros(3, dtype=pair)
    # element 0: value peaks at rank 2; element 1: all equal (tie -> lowest index); element 2: falls with rank
    mine["v"] = [-abs(rank - 2.0), 5.0, -float(rank)]
    mine["i"] = rank
    out = np.zeros(3, dtype=pair)
    mpiT.Allreduce(mine.view(np.uint8), out.view(np.uint8), 3, mpiT.FLOAT_INT, mpiT.MAXLOC)
    assert out["v"].tolist() == [0.0, 5.0, 0.0] and out["i"].tolist() == [2, 0, 0], (rank, out)
    mpiT.Reduce(mine.view(np.uint8), out.view(np.uint8), 3, mpiT.FLOAT_INT, mpiT.MINLOC, 1)
    if rank == 1:
        assert out["v"].tolist() == [-2.0, 5.0, -float(size - 1)] and out["i"].tolist() == [0, 0, size - 1], out
    two = torch.tensor([[rank * 10, rank], [7, rank]], dtype=torch.int32)
    res = torch.zeros_like(two)
    mpiT.Allreduce(two, res, 2, getattr(mpiT, "2INT"), mpiT.MAXLOC)
    assert res.tolist() == [[(size - 1) * 10, size - 1], [7, 0]]
    us = np.array([rank + 1, 60000], dtype=np.uint16)
    tot = np.zeros(2, dtype=np.uint16)
    mpiT.Allreduce(us, tot, 2, mpiT.UNSIGNED_SHORT, mpiT.MAX)
    assert tot.tolist() == [size, 60000]
    ul = np.array([2 ** 40 + rank], dtype=np.uint64)
    s = np.zeros(1, dtype=np.uint64)
    mpiT.Allreduce(ul, s, 1, mpiT.UNSIGNED_LONG, mpiT.SUM)
    assert int(s[0]) == size * 2 ** 40 + size * (size - 1) // 2
    ld = np.array([1.0 + rank], dtype=np.longdouble)
    p = np.zeros(1, dtype=np.longdouble)
    mpiT.Allreduce(ld.view(np.uint8), p.view(np.uint8), 1, mpiT.LONG_DOUBLE, mpiT.PROD)
    want = 1.0
    for r in range(size):
        want *= 1.0 + r
    assert float(p[0]) == want
    trip = mpiT.Type_commit(mpiT.Type_contiguous(3, mpiT.FLOAT))
    v = torch.full((6,), float(rank))
    w = torch.zeros(6)
    mpiT.Allreduce(v, w, 2, trip, mpiT.SUM)
    assert w.tolist() == [size * (size - 1) / 2.0] * 6
    try:
        mpiT.Allreduce(v, w, 3, mpiT.FLOAT, mpiT.MAXLOC)
        raise AssertionError("MAXLOC over a scalar type")
    except mpiT.MPIError as e:
        assert e.code == mpiT.ERR_OP


def case_c_topology():
    """Cartesian grid: coordinates, shifts with and without wrap-around, sub-grids; graph neighbours."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    assert size == 6
    dims = mpiT.Dims_create(size, [0, 0])
    assert dims == [3, 2]
    cart = mpiT.Cart_create(mpiT.COMM_WORLD, dims, [True, False])
    assert mpiT.Topo_test(cart) == mpiT.CART and mpiT.Cartdim_get(cart) == 2
    me = mpiT.Comm_rank(cart)
    coords = mpiT.Cart_coords(cart, me)
    assert coords == [me // 2, me % 2] and mpiT.Cart_rank(cart, coords) == me
    assert mpiT.Cart_get(cart) == ([3, 2], [True, False], coords)
    assert mpiT.Cart_rank(cart, [coords[0] + 3, coords[1]]) == me            # periodic dimension wraps
    src, dst = mpiT.Cart_shift(cart, 0, 1)
    assert dst == ((coords[0] + 1) % 3) * 2 + coords[1] and src == ((coords[0] - 1) % 3) * 2 + coords[1]
    s1, d1 = mpiT.Cart_shift(cart, 1, 1)
    assert d1 == (mpiT.PROC_NULL if coords[1] == 1 else me + 1) and s1 == (mpiT.PROC_NULL if coords[0 + 1] == 0 else me - 1)
    # halo exchange along the periodic dimension
    got = torch.zeros(1)
    mpiT.Sendrecv(torch.tensor([float(me)]), 1, mpiT.FLOAT, dst, 1, got, 1, mpiT.FLOAT, src, 1, cart)
    assert float(got) == src
    rows = mpiT.Cart_sub(cart, [False, True])                                  # 3 sub-grids of 2
    assert mpiT.Comm_size(rows) == 2 and mpiT.Comm_rank(rows) == coords[1] and mpiT.Cartdim_get(rows) == 1
    tot = torch.zeros(1)
    mpiT.Allreduce(torch.tensor([float(me)]), tot, 1, mpiT.FLOAT, mpiT.SUM, rows)
    assert float(tot) == 4 * coords[0] + 1
    assert mpiT.Cart_map(mpiT.COMM_WORLD, [2, 2], [False, False]) == (rank if rank < 4 else mpiT.UNDEFINED)
    small = mpiT.Cart_create(mpiT.COMM_WORLD, [2, 2], [False, False])
    assert (small == mpiT.COMM_NULL) == (rank >= 4)
    dup = mpiT.Comm_dup(cart)
    assert mpiT.Topo_test(dup) == mpiT.CART and mpiT.Cart_coords(dup, me) == coords
    # a ring as a graph
    index = [2 * (i + 1) for i in range(size)]
    edges = [e for i in range(size) for e in ((i - 1) % size, (i + 1) % size)]
    ring = mpiT.Graph_create(mpiT.COMM_WORLD, index, edges)
    assert mpiT.Topo_test(ring) == mpiT.GRAPH and mpiT.Graphdims_get(ring) == (size, 2 * size)
    assert mpiT.Graph_neighbors_count(ring, rank) == 2
    assert mpiT.Graph_neighbors(ring, rank) == [(rank - 1) % size, (rank + 1) % size]
    assert mpiT.Graph_get(ring) == (index, edges) and mpiT.Graph_map(mpiT.COMM_WORLD, index, edges) == rank
    assert mpiT.Topo_test(mpiT.COMM_WORLD) == mpiT.UNDEFINED
    mpiT.Comm_free(ring)
    assert mpiT.Topo_test(ring) == mpiT.UNDEFINED


def case_c_file_io():
    """Every rank writes its block of a 2-d array through a subarray view; ordered and shared-pointer appends."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    assert size == 4
    path = _shared_tmp("grid.bin")
    if rank == 0 and os.path.exists(path):
        os.unlink(path)
    mpiT.Barrier()
    fh = mpiT.File_open(mpiT.COMM_WORLD, path, mpiT.MODE_CREATE | mpiT.MODE_RDWR)
    G, B = 8, 4                                               # 8x8 floats, 2x2 process grid of 4x4 blocks
    r, c = divmod(rank, 2)
    view = mpiT.Type_commit(mpiT.Type_create_subarray([G, G], [B, B], [r * B, c * B], mpiT.ORDER_C, mpiT.FLOAT))
    mpiT.File_set_view(fh, 32, mpiT.FLOAT, view)             # 32-byte header
    full = torch.arange(G * G, dtype=torch.float32).view(G, G)
    mine = full[r * B:(r + 1) * B, c * B:(c + 1) * B].contiguous()
    st = mpiT.Status()
    mpiT.File_write_all(fh, mine, B * B, mpiT.FLOAT, st)
    assert mpiT.Get_count(st, mpiT.FLOAT) == B * B
    mpiT.File_sync(fh)
    # everyone reads the whole array back through a flat view
    mpiT.File_set_view(fh, 32, mpiT.FLOAT, mpiT.FLOAT)
    whole = torch.zeros(G, G)
    mpiT.File_read_at_all(fh, 0, whole, G * G, mpiT.FLOAT)
    assert torch.equal(whole, full)
    assert mpiT.File_get_size(fh) == 32 + G * G * 4
    # read my block into a strided region of a larger tensor with a memory datatype
    mpiT.File_set_view(fh, 32, mpiT.FLOAT, view)
    big = torch.zeros(G, G)
    mpiT.File_read_at(fh, 0, big, 1, view)                    # same shape in memory as in the file
    assert torch.equal(big[r * B:(r + 1) * B, c * B:(c + 1) * B], mine) and float(big.sum()) == float(mine.sum())
    # ordered append: rank-ordered records of different lengths after the array
    mpiT.File_set_view(fh, 32 + G * G * 4, mpiT.INT, mpiT.INT)
    rec = torch.full((rank + 1,), rank, dtype=torch.int32)
    mpiT.File_write_ordered(fh, rec, rank + 1, mpiT.INT)
    mpiT.File_write_ordered_begin(fh, rec + 10, rank + 1, mpiT.INT)
    mpiT.File_write_ordered_end(fh, rec)
    mpiT.Barrier()
    assert mpiT.File_get_position_shared(fh) == 20
    mpiT.Barrier()                                            # nobody appends before everyone has looked
    allrec = torch.zeros(20, dtype=torch.int32)
    mpiT.File_read_at(fh, 0, allrec, 20, mpiT.INT)
    want = [p for p in range(4) for _ in range(p + 1)]
    assert allrec.tolist() == want + [w + 10 for w in want]
    # shared pointer: every rank appends one record, all land, none overlap
    mpiT.File_write_shared(fh, torch.tensor([100 + rank], dtype=torch.int32), 1, mpiT.INT)
    mpiT.Barrier()
    tail = torch.zeros(4, dtype=torch.int32)
    mpiT.File_read_at(fh, 20, tail, 4, mpiT.INT)
    assert sorted(tail.tolist()) == [100, 101, 102, 103]
    mpiT.File_seek_shared(fh, 2, mpiT.SEEK_SET)
    assert mpiT.File_get_position_shared(fh) == 2
    g = mpiT.File_get_group(fh)
    assert mpiT.Group_size(g) == size
    mpiT.File_set_size(fh, 32)
    mpiT.Barrier()
    assert mpiT.File_get_size(fh) == 32
    mpiT.File_close(fh)
    # exclusive create: collective outcome, error on every rank
    try:
        mpiT.File_open(mpiT.COMM_WORLD, path, mpiT.MODE_CREATE | mpiT.MODE_EXCL | mpiT.MODE_WRONLY)
        raise AssertionError("exclusive create of an existing file succeeded")
    except mpiT.MPIError as e:
        assert e.code == mpiT.ERR_FILE_EXISTS
    fh = mpiT.File_open(mpiT.COMM_WORLD, path, mpiT.MODE_RDWR | mpiT.MODE_DELETE_ON_CLOSE)
    mpiT.File_close(fh)
    assert not os.path.exists(path)


def case_c_pscw_and_attrs():
    """Post / start / complete / wait epochs between pairs; Comm attributes across Comm_dup; Request_get_status."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    base = torch.zeros(8)
    win = mpiT.Win_create(base, 32, 4, None, mpiT.COMM_WORLD)
    world = mpiT.Win_get_group(win)
    assert world.ranks == list(range(size))
    # even ranks are targets of their right neighbour and of rank size-1; odd ranks are origins
    if rank % 2 == 0:
        origins = sorted({rank + 1, size - 1})
        mpiT.Win_post(mpiT.Group_incl(world, origins), 0, win)
        while not mpiT.Win_test(win):
            pass
        want = torch.zeros(8)
        want[0:2] = float(rank + 1)
        want[4] = sum(float(o) for o in origins)
        assert torch.equal(base, want), (rank, base)
    else:
        targets = [rank - 1] if rank != size - 1 else list(range(0, size, 2))
        mpiT.Win_start(mpiT.Group_incl(world, targets), 0, win)
        got = {}
        for t in targets:
            if t == rank - 1:
                mpiT.Put(torch.full((2,), float(rank)), 2, mpiT.FLOAT, t, 0, win)
            mpiT.Accumulate(torch.tensor([float(rank)]), 1, mpiT.FLOAT, t, 4, mpiT.SUM, win)
            got[t] = torch.full((1,), -1.0)
            mpiT.Get(got[t], 1, mpiT.FLOAT, t, 7, win)
        mpiT.Win_complete(win)
        assert all(float(v) == 0.0 for v in got.values())
        try:
            mpiT.Win_complete(win)
            raise AssertionError("Win_complete without an access epoch")
        except mpiT.MPIError as e:
            assert e.code == mpiT.ERR_RMA_SYNC
    mpiT.Barrier()
    mpiT.Win_free(win)
    # attributes follow Comm_dup through the copy callback and die with Comm_free
    freed = []
    kv = mpiT.Comm_create_keyval(lambda old, k, state, v: (v + 1, True), lambda c, k, v, s: freed.append(v))
    mpiT.Comm_set_attr(mpiT.COMM_WORLD, kv, 41)
    dup = mpiT.Comm_dup(mpiT.COMM_WORLD)
    assert mpiT.Comm_get_attr(dup, kv) == (42, True) and mpiT.Comm_get_attr(mpiT.COMM_WORLD, kv) == (41, True)
    assert mpiT.Comm_get_attr(mpiT.COMM_WORLD, mpiT.TAG_UB)[0] >= 32767
    assert mpiT.Comm_get_attr(mpiT.COMM_WORLD, mpiT.UNIVERSE_SIZE) == (size, True)
    mpiT.Comm_free(dup)
    assert freed == [42]
    mpiT.Comm_delete_attr(mpiT.COMM_WORLD, kv)
    assert freed == [42, 41] and mpiT.Comm_get_attr(mpiT.COMM_WORLD, kv) == (None, False)
    # Request_get_status does not release the request
    buf = torch.zeros(2)
    req = mpiT.Request(1)
    mpiT.Irecv(buf, 2, mpiT.FLOAT, (rank - 1) % size, 4, mpiT.COMM_WORLD, req)
    mpiT.Barrier()
    mpiT.Send(torch.tensor([1.0, float(rank)]), 2, mpiT.FLOAT, (rank + 1) % size, 4)
    st = mpiT.Status()
    while not mpiT.Request_get_status(req, st):
        pass
    assert st.source == (rank - 1) % size and st.tag == 4 and not req.is_null()
    assert mpiT.Request_get_status(req)                                         # still there
    mpiT.Wait(req)
    assert req.is_null() and buf.tolist() == [1.0, float((rank - 1) % size)]


def case_c_intercomm():
    """Servers (first half) and clients (second half) as an inter-communicator: create, address the remote group,
    merge; then the same pairing through a published port with accept / connect; Comm_join over a socket pair."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    assert size == 4
    half = size // 2
    server = rank < half
    local = mpiT.Comm_split(mpiT.COMM_WORLD, 0 if server else 1, rank)
    inter = mpiT.Intercomm_create(local, 0, mpiT.COMM_WORLD, half if server else 0, 77)
    assert mpiT.Comm_test_inter(inter) and not mpiT.Comm_test_inter(local)
    assert mpiT.Comm_remote_size(inter) == half and mpiT.Comm_size(inter) == half
    assert mpiT.Comm_remote_group(inter).ranks == (list(range(half, size)) if server else list(range(half)))
    lr = mpiT.Comm_rank(inter)
    # client i sends to server i (remote rank i); the server answers; sources are remote-group ranks
    st = mpiT.Status()
    if server:
        got = torch.zeros(1)
        mpiT.Recv(got, 1, mpiT.FLOAT, mpiT.ANY_SOURCE, 1, inter, st)
        assert st.source == lr and float(got) == 100 + half + lr
        mpiT.Send(got + 1, 1, mpiT.FLOAT, st.source, 2, inter)
    else:
        mpiT.Send(torch.tensor([100.0 + rank]), 1, mpiT.FLOAT, lr, 1, inter)
        back = torch.zeros(1)
        req = mpiT.Request(1)
        mpiT.Irecv(back, 1, mpiT.FLOAT, lr, 2, inter, req)
        mpiT.Wait(req, st)
        assert st.source == lr and float(back) == 101 + rank
    # merge with the clients first
    merged = mpiT.Intercomm_merge(inter, high=server)
    assert mpiT.Comm_size(merged) == size
    assert mpiT.Comm_rank(merged) == (half + lr if server else lr)
    tot = torch.zeros(1)
    mpiT.Allreduce(torch.ones(1), tot, 1, mpiT.FLOAT, mpiT.SUM, merged)
    assert float(tot) == size
    mpiT.Comm_disconnect(inter)
    # the same through a port published under a service name
    os.environ["MPIT_NAMESERVICE_DIR"] = _shared_tmp("names")
    if server:
        port = None
        if lr == 0:
            port = mpiT.Open_port()
            mpiT.Publish_name("param-server", None, port)
        mpiT.Barrier()                                   # the name is published before anybody looks it up
        inter2 = mpiT.Comm_accept(port or "", None, 0, local) if lr == 0 else \
            mpiT.Comm_accept("mpit-port:0:1", None, 0, local)
    else:
        mpiT.Barrier()
        port = mpiT.Lookup_name("param-server")
        inter2 = mpiT.Comm_connect(port, None, 0, local)
    assert mpiT.Comm_remote_size(inter2) == half
    x = torch.tensor([float(rank)])
    y = torch.zeros(1)
    mpiT.Sendrecv(x, 1, mpiT.FLOAT, lr, 3, y, 1, mpiT.FLOAT, lr, 3, inter2)
    assert float(y) == (rank + half if server else rank - half)
    mpiT.Barrier()
    if server and lr == 0:
        mpiT.Unpublish_name("param-server", None, port)
        mpiT.Close_port(port)
        try:
            mpiT.Lookup_name("param-server")
            raise AssertionError("lookup of an unpublished name succeeded")
        except mpiT.MPIError as e:
            assert e.code == mpiT.ERR_NAME
    assert mpiT.Comm_get_parent() == mpiT.COMM_NULL
    try:
        mpiT.Comm_spawn("worker", None, 2)
        raise AssertionError("spawn")
    except mpiT.MPIError as e:
        assert e.code == mpiT.ERR_SPAWN
    mpiT.Comm_free(inter2)
    # Comm_join: ranks 0 and 1 connect a TCP socket and join over it
    if rank in (0, 1):
        addr_file = _shared_tmp("join-port")
        if rank == 0:
            srv = socket.socket()
            srv.bind(("127.0.0.1", 0))
            srv.listen(1)
            with open(addr_file + ".tmp", "w") as f:
                f.write(str(srv.getsockname()[1]))
            os.replace(addr_file + ".tmp", addr_file)
            conn, _ = srv.accept()
        else:
            import time
            while not os.path.exists(addr_file):
                time.sleep(0.01)
            conn = socket.create_connection(("127.0.0.1", int(open(addr_file).read())))
        pair = mpiT.Comm_join(conn.fileno())
        assert mpiT.Comm_remote_size(pair) == 1 and mpiT.Comm_remote_group(pair).ranks == [1 - rank]
        z = torch.zeros(1)
        mpiT.Sendrecv(torch.tensor([float(rank)]), 1, mpiT.FLOAT, 0, 8, z, 1, mpiT.FLOAT, 0, 8, pair)
        assert float(z) == 1 - rank
        conn.close()
        if rank == 0:
            os.unlink(addr_file)




def case_c_rendezvous():
    """Large non-blocking sends between processes of one host: single-copy rendezvous (process_vm_readv).  Both
    arrival orders, wildcards, probe counts, truncation, ordering against small eager messages, cancellation of an
    unmatched send.  With MPIT_CMA=0 the same body runs over the socket path."""
    import time
    rank, size = mpiT.get_rank(), mpiT.get_size()
    n = 1 << 20                                           # 4 MiB of floats: far above the 256 KiB threshold
    if os.environ.get("MPIT_CMA", "1") == "0":
        assert mpiT.stats()["rendezvous_peers"] == 0
    cma = mpiT.stats()["rendezvous_peers"] == size - 1    # false where the kernel forbids process_vm_readv
    right, left = (rank + 1) % size, (rank - 1) % size
    src = torch.arange(n, dtype=torch.float32) + rank * 1e6
    # 1. receive posted first (descriptor meets a posted receive on the progress thread)
    dst = torch.zeros(n)
    req = mpiT.Request(2)
    mpiT.Irecv(dst, n, mpiT.FLOAT, left, 11, mpiT.COMM_WORLD, req, 0)
    mpiT.Barrier()
    mpiT.Isend(src, n, mpiT.FLOAT, right, 11, mpiT.COMM_WORLD, req, 1)
    st = mpiT.Status()
    mpiT.Wait(req, st, 0)
    mpiT.Wait(req, None, 1)
    assert st.source == left and st.tag == 11 and mpiT.Get_count(st, mpiT.FLOAT) == n
    assert torch.equal(dst, torch.arange(n, dtype=torch.float32) + left * 1e6)
    # 2. send first: the message is visible to Iprobe with its full size while the data is still with the sender
    before = mpiT.stats()["bytes_recv"]
    mpiT.Isend(src, n, mpiT.FLOAT, right, 12, mpiT.COMM_WORLD, req, 1)
    while not mpiT.Iprobe(mpiT.ANY_SOURCE, 12, mpiT.COMM_WORLD, st):
        time.sleep(0.001)
    assert st.source == left and mpiT.Get_count(st, mpiT.FLOAT) == n
    if cma:
        assert not mpiT.Test(req, None, 1)                # not complete before the receiver has pulled it
        assert mpiT.stats()["bytes_recv"] - before < n    # ... and nothing was buffered on this side
    mpiT.Barrier()                                        # nobody pulls before everybody has looked
    dst.zero_()
    mpiT.Recv(dst, n, mpiT.FLOAT, mpiT.ANY_SOURCE, 12, mpiT.COMM_WORLD, st)
    assert st.source == left and torch.equal(dst, torch.arange(n, dtype=torch.float32) + left * 1e6)
    mpiT.Wait(req, None, 1)
    # 3. non-overtaking: small eager, large rendezvous, small eager on one tag arrive in that order
    a, c = torch.tensor([1.0 + rank]), torch.tensor([3.0 + rank])
    r3 = mpiT.Request(3)
    mpiT.Isend(a, 1, mpiT.FLOAT, right, 13, mpiT.COMM_WORLD, r3, 0)
    mpiT.Isend(src, n, mpiT.FLOAT, right, 13, mpiT.COMM_WORLD, r3, 1)
    mpiT.Isend(c, 1, mpiT.FLOAT, right, 13, mpiT.COMM_WORLD, r3, 2)
    x, z = torch.zeros(1), torch.zeros(1)
    dst.zero_()
    mpiT.Recv(x, 1, mpiT.FLOAT, left, 13)
    mpiT.Recv(dst, n, mpiT.FLOAT, left, 13)
    mpiT.Recv(z, 1, mpiT.FLOAT, left, 13)
    assert float(x) == 1.0 + left and float(z) == 3.0 + left and float(dst[5]) == 5 + left * 1e6
    mpiT.Waitall(r3)
    # 4. truncation: a short receive gets the head of the message and the error
    mpiT.Isend(src, n, mpiT.FLOAT, right, 14, mpiT.COMM_WORLD, req, 1)
    small = torch.zeros(100)
    try:
        mpiT.Recv(small, 100, mpiT.FLOAT, left, 14, mpiT.COMM_WORLD, st)
        raise AssertionError("truncation not reported")
    except mpiT.MPIError as e:
        assert e.code == mpiT.ERR_TRUNCATE
    assert torch.equal(small, torch.arange(100, dtype=torch.float32) + left * 1e6)
    mpiT.Wait(req, None, 1)                               # the sender still completes
    # 5. Sendrecv of large buffers in both directions at once
    back = torch.zeros(n)
    mpiT.Sendrecv(src, n, mpiT.FLOAT, right, 15, back, n, mpiT.FLOAT, left, 15)
    assert float(back[-1]) == n - 1 + left * 1e6
    # 6. cancelling a large send: under the rendezvous only the descriptor has left, and the receiver drops it if no
    #    receive has taken it (the send completes as cancelled, nothing arrives); over the socket the payload is
    #    already travelling, the cancel comes too late and the message is delivered
    mpiT.Barrier()
    if rank == 0:
        mpiT.Isend(src, n, mpiT.FLOAT, 1, 16, mpiT.COMM_WORLD, req, 1)
        time.sleep(0.05)
        mpiT.Cancel(req, 1)
        mpiT.Wait(req, st, 1)
        cancelled = mpiT.Test_cancelled(st)
        assert cancelled == cma
        mpiT.Send(torch.tensor([1.0 if cancelled else 0.0]), 1, mpiT.FLOAT, 1, 17)
    elif rank == 1:
        flag = torch.zeros(1)
        mpiT.Recv(flag, 1, mpiT.FLOAT, 0, 17)
        if float(flag):
            assert not mpiT.Iprobe(0, 16)
        else:
            dst.zero_()
            mpiT.Recv(dst, n, mpiT.FLOAT, 0, 16)
            assert float(dst[7]) == 7.0
    mpiT.Barrier()
    # 7. ... and a cancel that arrives after the receive has matched changes nothing
    if rank == 0:
        mpiT.Isend(src, n, mpiT.FLOAT, 1, 18, mpiT.COMM_WORLD, req, 1)
        mpiT.Recv(None, 0, mpiT.BYTE, 1, 19)              # rank 1 has received it
        mpiT.Cancel(req, 1)
        mpiT.Wait(req, st, 1)
        assert not mpiT.Test_cancelled(st)
    elif rank == 1:
        dst.zero_()
        mpiT.Recv(dst, n, mpiT.FLOAT, 0, 18)
        assert float(dst[9]) == 9.0
        mpiT.Send(None, 0, mpiT.BYTE, 0, 19)
    mpiT.Barrier()


def case_c_sharded_checkpoint():
    """Every rank writes its shard of a flat vector into one Torch7 file through MPI-IO; the file is what torch.save of
    the whole tensor would have produced; shards read back in a different partition."""
    from mpit_b200 import t7
    from mpit_b200.parallel.center import partition
    from mpit_b200.utils.sharded_ckpt import load_flat_sharded, save_flat_sharded
    rank, size = mpiT.get_rank(), mpiT.get_size()
    n = 351946                                             # the CIFAR CNN's parameter count: uneven shards
    full = torch.arange(n, dtype=torch.float32) * 0.5
    off, cnt = partition(n, size)[rank]
    path = _shared_tmp("centre.t7")
    save_flat_sharded(path, full[off:off + cnt].contiguous(), off, n)
    mpiT.Barrier()
    if rank == 0:
        whole = t7.load(path)
        assert whole.dtype == torch.float32 and torch.equal(whole, full)
        assert open(path, "rb").read() == t7.dumps(full)   # byte for byte the gathered save
    # read back with the shards rotated by one rank
    off2, cnt2 = partition(n, size)[(rank + 1) % size]
    piece = torch.zeros(cnt2)
    assert load_flat_sharded(path, piece, off2) == n
    assert torch.equal(piece, full[off2:off2 + cnt2])
    # a shorter vector over the same path leaves no tail behind
    mpiT.Barrier()
    m = 1000
    o3, c3 = partition(m, size)[rank]
    save_flat_sharded(path, full[o3:o3 + c3].contiguous(), o3, m)
    mpiT.Barrier()
    assert torch.equal(t7.load(path), full[:m])
    mpiT.Barrier()
    if rank == 0:
        os.unlink(path)


def case_c_ring_allreduce():
    """Large vectors on >= 3 ranks take the ring reduce-scatter / all-gather path: uneven chunking, several types and
    operations, in place, and the non-blocking form; a small vector still takes the tree."""
    rank, size = mpiT.get_rank(), mpiT.get_size()
    n = 300007                                             # not a multiple of the rank count; 1.2 MB of floats
    i = torch.arange(n)
    a = ((i % 97) * (rank + 1)).float()
    out = torch.zeros(n)
    mpiT.Allreduce(a, out, n, mpiT.FLOAT, mpiT.SUM)
    assert torch.equal(out, ((i % 97) * (size * (size + 1) // 2)).float())
    l = (i + rank * 7).to(torch.int64)
    lo, hi = torch.zeros(n, dtype=torch.int64), torch.zeros(n, dtype=torch.int64)
    mpiT.Allreduce(l, hi, n, mpiT.LONG, mpiT.MAX)
    mpiT.Allreduce(l, lo, n, mpiT.LONG, mpiT.MIN)
    assert torch.equal(hi, i + (size - 1) * 7) and torch.equal(lo, i)
    d = torch.full((n,), 1.0 + rank, dtype=torch.float64)
    mpiT.Allreduce(d, d, n, mpiT.DOUBLE, mpiT.PROD)         # in place
    want = 1.0
    for r in range(size):
        want *= 1.0 + r
    assert float(d[0]) == want and float(d[-1]) == want and float(d.min()) == float(d.max())
    bits = torch.full((n,), 1 << rank, dtype=torch.int32)
    orr = torch.zeros(n, dtype=torch.int32)
    mpiT.Allreduce(bits, orr, n, mpiT.INT, mpiT.BOR)
    assert int(orr[0]) == (1 << size) - 1 and int(orr[n // 2]) == (1 << size) - 1
    req = mpiT.Request(1)
    out.zero_()
    mpiT.Iallreduce(a, out, n, mpiT.FLOAT, mpiT.SUM, mpiT.COMM_WORLD, req)
    mpiT.Wait(req)
    assert torch.equal(out, ((i % 97) * (size * (size + 1) // 2)).float())
    # sub-communicator of three out of the ranks, and a small vector (tree path) for contrast
    sub = mpiT.Comm_split(mpiT.COMM_WORLD, 0 if rank < 3 else 1, rank)
    if rank < 3:
        s3 = torch.zeros(n)
        mpiT.Allreduce(a, s3, n, mpiT.FLOAT, mpiT.SUM, sub)
        assert torch.equal(s3, ((i % 97) * 6).float())
    tiny = torch.zeros(5)
    mpiT.Allreduce(torch.full((5,), float(rank)), tiny, 5, mpiT.FLOAT, mpiT.SUM)
    assert tiny.tolist() == [size * (size - 1) / 2.0] * 5
    try:
        mpiT.Allreduce(a, out, n, mpiT.FLOAT, mpiT.BOR)
        raise AssertionError("bitwise or of floats")
    except mpiT.MPIError as e:
        assert e.code == mpiT.ERR_OP


def case_c_big_collectives():
    """Alltoall / Gather / Scatter / Scan / Allgather with 4 MiB blocks: the sizes at which their sends go through the
    single-copy rendezvous (and through the socket with MPIT_CMA=0)."""
    r, n = mpiT.get_rank(), mpiT.get_size()
    c = 1 << 20   # 4 MB per rank pair
    idx = torch.arange(n * c)
    s = (idx % 1000 + (idx // c) * 10000 + r * 100000).float()      # encodes (sender, destination block, position)
    out = torch.zeros(n * c)
    mpiT.Alltoall(s, c, mpiT.FLOAT, out)
    for p in range(n):
        blk = out[p * c:(p + 1) * c]
        assert torch.equal(blk, (torch.arange(r * c, (r + 1) * c) % 1000 + r * 10000 + p * 100000).float()), (r, p)
    g = torch.zeros(n * c) if r == 1 else None
    mpiT.Gather(s[:c], c, mpiT.FLOAT, g, 1)
    if r == 1:
        for p in range(n):
            assert torch.equal(g[p * c:(p + 1) * c], (torch.arange(c) % 1000 + p * 100000).float())
    piece = torch.zeros(c)
    mpiT.Scatter(s if r == 2 else None, c, mpiT.FLOAT, piece, 2)
    assert torch.equal(piece, (torch.arange(r * c, (r + 1) * c) % 1000 + r * 10000 + 200000).float())
    sc = torch.zeros(c)
    mpiT.Scan(torch.full((c,), 1.0 + r), sc, c, mpiT.FLOAT, mpiT.SUM)
    assert float(sc[9]) == sum(1.0 + q for q in range(r + 1)) and float(sc.min()) == float(sc.max())
    ag = torch.zeros(n * c)
    mpiT.Allgather(s[:c], c, mpiT.FLOAT, ag)
    for p in range(n):
        assert torch.equal(ag[p * c:(p + 1) * c], (torch.arange(c) % 1000 + p * 100000).float())
    # many large non-blocking sends and receives in flight at once (Alltoallv posts them all, then waits)
    counts = [c // 2 + 1000 * q for q in range(n)]              # what I send to q; I receive c // 2 + 1000 * r from each
    sd = [sum(counts[:q]) for q in range(n)]
    mine = c // 2 + 1000 * r
    sv = torch.cat([torch.full((counts[q],), float(r * 10 + q)) for q in range(n)])
    rv = torch.zeros(n * mine)
    mpiT.Alltoallv(sv, counts, sd, mpiT.FLOAT, rv, [mine] * n, [q * mine for q in range(n)])
    for q in range(n):
        seg = rv[q * mine:(q + 1) * mine]
        assert float(seg.min()) == float(seg.max()) == q * 10 + r, (r, q)


CASES = {k[5:]: v for k, v in globals().items() if k.startswith("case_")}
