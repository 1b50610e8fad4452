"""Single-process checks of the MPI chapters beyond point-to-point and collectives: derived datatypes (against
NumPy slicing as the oracle), MPI-IO on COMM_SELF files, Info, keyvals / attributes, error handlers, Dims_create,
generalized requests; plus the coverage of the reference's 270 spec files.  The multi-process side is in
``mp_cases_chapters.py`` (run from test_msg_layer.py)."""
import os
import threading

import numpy as np
import pytest
import torch

from mpit_b200 import msg as mpiT


# ---- derived datatypes ---------------------------------------------------------------------------------------------
def _packed(buf, count, dtype):
    out = np.zeros(mpiT.Pack_size(count, dtype), dtype=np.uint8)
    end = mpiT.Pack(buf, count, dtype, out, 0)
    assert end == out.size
    return out


def test_vector_is_a_strided_column():
    a = np.arange(6 * 8, dtype=np.float32).reshape(6, 8)
    col = mpiT.Type_commit(mpiT.Type_vector(6, 1, 8, mpiT.FLOAT))
    assert mpiT.Type_size(col) == 24 and mpiT.Type_get_extent(col) == (0, (5 * 8 + 1) * 4)
    got = _packed(a.reshape(-1)[3:], 1, col).view(np.float32)
    assert np.array_equal(got, a[:, 3])
    # two columns in a row need an element extent of one float: resize
    col1 = mpiT.Type_commit(mpiT.Type_create_resized(col, 0, 4))
    got = _packed(a, 2, col1).view(np.float32)
    assert np.array_equal(got, np.concatenate([a[:, 0], a[:, 1]]))
    # scatter back into a zeroed array
    b = np.zeros_like(a)
    mpiT.Unpack(got.view(np.uint8), 0, b, 2, col1)
    want = np.zeros_like(a)
    want[:, :2] = a[:, :2]
    assert np.array_equal(b, want)
    assert mpiT.Type_get_envelope(col)[3] == mpiT.COMBINER_VECTOR
    assert mpiT.Type_get_contents(col) == ([6, 1, 8], [], [mpiT.FLOAT])
    assert mpiT.Type_free(col) == mpiT.DATATYPE_NULL


def test_indexed_hindexed_block_and_contiguous():
    x = torch.arange(40, dtype=torch.float64)
    idx = mpiT.Type_commit(mpiT.Type_indexed([2, 3, 1], [10, 0, 30], mpiT.DOUBLE))
    got = torch.from_numpy(_packed(x, 1, idx).view(np.float64))
    assert torch.equal(got, torch.cat([x[10:12], x[0:3], x[30:31]]))          # type-map order, not address order
    assert mpiT.Type_get_true_extent(idx) == (0, 31 * 8)
    hidx = mpiT.Type_commit(mpiT.Type_create_hindexed([2, 3, 1], [80, 0, 240], mpiT.DOUBLE))
    assert np.array_equal(_packed(x, 1, hidx), _packed(x, 1, idx))
    blk = mpiT.Type_commit(mpiT.Type_create_indexed_block(2, [4, 0], mpiT.DOUBLE))
    assert torch.equal(torch.from_numpy(_packed(x, 1, blk).view(np.float64)), torch.cat([x[4:6], x[0:2]]))
    pair = mpiT.Type_commit(mpiT.Type_contiguous(2, blk))
    assert mpiT.Type_size(pair) == 64
    got = torch.from_numpy(_packed(x, 1, pair).view(np.float64))
    assert torch.equal(got, torch.cat([x[4:6], x[0:2], x[10:12], x[6:8]]))   # second copy one extent (6 doubles) on
    hv = mpiT.Type_commit(mpiT.Type_create_hvector(3, 2, 40, mpiT.DOUBLE))
    assert torch.equal(torch.from_numpy(_packed(x, 1, hv).view(np.float64)), torch.cat([x[0:2], x[5:7], x[10:12]]))
    d = mpiT.Type_dup(hv)
    assert mpiT.Type_get_envelope(d)[3] == mpiT.COMBINER_DUP and mpiT.Type_size(d) == mpiT.Type_size(hv)


def test_struct_with_mixed_fields_and_get_address():
    rec = np.zeros(3, dtype=np.dtype([("id", np.int32), ("w", np.float64, (2,)), ("tag", np.int8)], align=True))
    rec["id"], rec["w"], rec["tag"] = [1, 2, 3], [[.5, 1.5], [2.5, 3.5], [4.5, 5.5]], [7, 8, 9]
    base = mpiT.Get_address(rec)
    offs = [rec.dtype.fields[k][1] for k in ("id", "w", "tag")]
    st = mpiT.Type_commit(mpiT.Type_create_struct([1, 2, 1], offs, [mpiT.INT, mpiT.DOUBLE, mpiT.CHAR]))
    assert base == rec.ctypes.data
    assert mpiT.Type_get_extent(st) == (0, rec.dtype.itemsize)               # trailing padding like the C struct
    assert mpiT.Type_size(st) == 4 + 16 + 1
    raw = rec.view(np.uint8).reshape(-1)
    stream = _packed(raw, 3, st)
    back = np.zeros_like(raw)
    mpiT.Unpack(stream, 0, back, 3, st)
    again = back.view(rec.dtype)
    assert np.array_equal(again["id"], rec["id"]) and np.array_equal(again["w"], rec["w"]) \
        and np.array_equal(again["tag"], rec["tag"])
    # external32: fixed big-endian layout, round trip
    n = mpiT.Pack_external_size("external32", 3, st)
    assert n == 3 * 21
    ext = np.zeros(n, dtype=np.uint8)
    assert mpiT.Pack_external("external32", raw, 3, st, ext, 0) == n
    assert ext[:4].tolist() == [0, 0, 0, 1]                                   # id = 1, big endian
    back2 = np.zeros_like(raw)
    mpiT.Unpack_external("external32", ext, 0, back2, 3, st)
    assert np.array_equal(back2.view(rec.dtype)["w"], rec["w"])
    with pytest.raises(mpiT.MPIError):
        mpiT.Pack_external("ieee-le", raw, 1, st, ext, 0)


@pytest.mark.parametrize("order", ["C", "F"])
def test_subarray_matches_numpy_slicing(order):
    sizes, sub, starts = (5, 6, 4), (2, 3, 2), (1, 2, 1)
    a = np.arange(np.prod(sizes), dtype=np.int32).reshape(sizes, order=order)
    t = mpiT.Type_commit(mpiT.Type_create_subarray(sizes, sub, starts,
                                                   mpiT.ORDER_C if order == "C" else mpiT.ORDER_FORTRAN, mpiT.INT))
    flat = a.reshape(-1, order=order).copy()
    got = _packed(flat, 1, t).view(np.int32)
    sl = a[1:3, 2:5, 1:3]
    assert np.array_equal(got, sl.reshape(-1, order=order))
    assert mpiT.Type_get_extent(t) == (0, a.size * 4)


def test_darray_pieces_partition_the_global_array():
    g = (6, 8)
    a = np.arange(48, dtype=np.float32).reshape(g)
    seen = np.zeros(48, dtype=np.int32)
    for rank in range(4):
        t = mpiT.Type_commit(mpiT.Type_create_darray(4, rank, g, [mpiT.DISTRIBUTE_BLOCK, mpiT.DISTRIBUTE_CYCLIC],
                                                     [mpiT.DISTRIBUTE_DFLT_DARG, 2], [2, 2], mpiT.ORDER_C, mpiT.FLOAT))
        got = _packed(a, 1, t).view(np.float32)
        r, c = divmod(rank, 2)
        cols = [j for j in range(8) if (j // 2) % 2 == c]
        assert np.array_equal(got, a[3 * r:3 * r + 3][:, cols].reshape(-1))
        seen[got.astype(np.int64)] += 1
    assert (seen == 1).all()


def test_datatype_queries_and_errors():
    assert mpiT.Type_match_size(mpiT.TYPECLASS_REAL, 8) == mpiT.DOUBLE
    assert mpiT.Type_match_size(mpiT.TYPECLASS_INTEGER, 4) == mpiT.INT
    t = mpiT.Type_vector(2, 1, 3, mpiT.FLOAT)
    with pytest.raises(mpiT.MPIError):                      # not committed
        mpiT.Pack(np.zeros(8, np.float32), 1, t, np.zeros(8, np.uint8), 0)
    mpiT.Type_commit(t)
    with pytest.raises(mpiT.MPIError) as e:                 # buffer too short for the span
        mpiT.Pack(np.zeros(3, np.float32), 1, t, np.zeros(8, np.uint8), 0)
    assert e.value.code == mpiT.ERR_TRUNCATE
    mpiT.Type_set_name(t, "two floats, stride 3")
    assert mpiT.Type_get_name(t) == "two floats, stride 3" and mpiT.Type_get_name(mpiT.FLOAT) == "MPI_FLOAT"
    st = mpiT.Status()
    mpiT.Status_set_elements(st, t, 3)                      # one and a half elements
    assert st.count_bytes == 12 and mpiT.Get_elements(st, t) == 3 and mpiT.Get_count(st, t) == mpiT.UNDEFINED
    mpiT.Status_set_cancelled(st, True)
    assert mpiT.Test_cancelled(st)
    with pytest.raises(mpiT.MPIError):
        mpiT.Type_free(mpiT.FLOAT)


def test_remaining_predefined_types_and_bound_markers():
    """lua-mpi.h:130-151 lists 22 datatypes; the twelve without a native wire id are packed types here."""
    assert mpiT.Type_size(mpiT.UNSIGNED_SHORT) == 2 and mpiT.Type_size(mpiT.UNSIGNED_LONG) == 8
    assert mpiT.Type_size(mpiT.LONG_DOUBLE) == np.dtype(np.longdouble).itemsize and mpiT.LONG_LONG_INT == mpiT.LONG_LONG
    assert mpiT.Type_get_extent(mpiT.FLOAT_INT) == (0, 8) and mpiT.Type_get_extent(mpiT.DOUBLE_INT) == (0, 16)
    assert mpiT.Type_size(mpiT.DOUBLE_INT) == 12 and mpiT.Type_size(mpiT.SHORT_INT) == 6
    assert mpiT.Type_get_extent(mpiT.SHORT_INT) == (0, 8) and mpiT.Type_get_name(getattr(mpiT, "2INT")) == "MPI_2INT"
    assert mpiT.Type_get_envelope(mpiT.LONG_INT)[3] == mpiT.COMBINER_NAMED
    with pytest.raises(mpiT.MPIError):
        mpiT.Type_free(mpiT.PACKED)
    # a C array of {double, int} records packs to 12 bytes per record
    rec = np.zeros(3, dtype=np.dtype([("v", np.float64), ("i", np.int32)], align=True))
    rec["v"], rec["i"] = [1.5, -2.0, 7.25], [4, 5, 6]
    stream = _packed(rec.view(np.uint8).reshape(-1), 3, mpiT.DOUBLE_INT)
    assert stream.size == 36 and stream[:8].view(np.float64)[0] == 1.5 and stream[8:12].view(np.int32)[0] == 4
    # MPI-1 style bounds: {LB at -4, one float at 0, UB at 12}
    t = mpiT.Type_commit(mpiT.Type_create_struct([1, 1, 1], [-4, 0, 12], [mpiT.LB, mpiT.FLOAT, mpiT.UB]))
    assert mpiT.Type_get_extent(t) == (-4, 16) and mpiT.Type_get_true_extent(t) == (0, 4) and mpiT.Type_size(t) == 4
    us = np.array([1, 65535, 7], dtype=np.uint16)
    out = np.zeros(3, dtype=np.uint16)
    mpiT.Unpack(_packed(us, 3, mpiT.UNSIGNED_SHORT), 0, out, 3, mpiT.UNSIGNED_SHORT)
    assert np.array_equal(out, us)
    ext = np.zeros(mpiT.Pack_external_size("external32", 3, mpiT.UNSIGNED_SHORT), dtype=np.uint8)
    mpiT.Pack_external("external32", us, 3, mpiT.UNSIGNED_SHORT, ext, 0)
    assert ext.tolist() == [0, 1, 255, 255, 0, 7]


# ---- Info, attributes, error handlers, misc -----------------------------------------------------------------------
def test_info_object():
    info = mpiT.Info_create()
    mpiT.Info_set(info, "striping_factor", "8")
    mpiT.Info_set(info, "access_style", "write_once")
    assert mpiT.Info_get_nkeys(info) == 2 and mpiT.Info_get_nthkey(info, 1) == "access_style"
    assert mpiT.Info_get(info, "striping_factor") == ("8", True) and mpiT.Info_get(info, "nope") == (None, False)
    assert mpiT.Info_get_valuelen(info, "access_style") == (10, True)
    dup = mpiT.Info_dup(info)
    mpiT.Info_delete(info, "striping_factor")
    assert mpiT.Info_get_nkeys(info) == 1 and mpiT.Info_get_nkeys(dup) == 2
    with pytest.raises(mpiT.MPIError) as e:
        mpiT.Info_delete(info, "striping_factor")
    assert e.value.code == mpiT.ERR_INFO_NOKEY
    assert mpiT.Info_free(dup) is mpiT.INFO_NULL


def test_type_and_win_attributes_with_callbacks():
    deleted = []
    kv = mpiT.Type_create_keyval(mpiT.DUP_FN, lambda obj, k, v, s: deleted.append((obj, v)), "state")
    nocopy = mpiT.Type_create_keyval()
    t = mpiT.Type_contiguous(3, mpiT.INT)
    mpiT.Type_set_attr(t, kv, {"layout": "rgb"})
    mpiT.Type_set_attr(t, nocopy, 5)
    d = mpiT.Type_dup(t)
    assert mpiT.Type_get_attr(d, kv) == ({"layout": "rgb"}, True)
    assert mpiT.Type_get_attr(d, nocopy) == (None, False)            # NULL_COPY_FN: not propagated
    mpiT.Type_delete_attr(t, kv)
    assert deleted == [(t, {"layout": "rgb"})] and mpiT.Type_get_attr(t, kv) == (None, False)
    mpiT.Type_free(d)                                                # runs the delete callback of the copy
    assert len(deleted) == 2
    assert mpiT.Type_free_keyval(kv) == mpiT.KEYVAL_INVALID
    with pytest.raises(mpiT.MPIError):
        mpiT.Type_set_attr(t, kv, 1)
    win = mpiT.Win_create(torch.zeros(16), 64, 4, None, mpiT.COMM_SELF)
    assert mpiT.Win_get_attr(win, mpiT.WIN_SIZE) == (64, True) and mpiT.Win_get_attr(win, mpiT.WIN_DISP_UNIT) == (4, True)
    wk = mpiT.Win_create_keyval()
    mpiT.Win_set_attr(win, wk, "mine")
    assert mpiT.Win_get_attr(win, wk) == ("mine", True)
    mpiT.Win_set_name(win, "centre")
    assert mpiT.Win_get_name(win) == "centre"


def test_error_handlers_and_user_error_codes():
    calls = []
    eh = mpiT.Comm_create_errhandler(lambda comm, code: calls.append((comm, code)))
    assert mpiT.Comm_get_errhandler(mpiT.COMM_SELF) == mpiT.ERRORS_ARE_FATAL
    mpiT.Comm_set_errhandler(mpiT.COMM_SELF, eh)
    cls = mpiT.Add_error_class()
    code = mpiT.Add_error_code(cls)
    mpiT.Add_error_string(code, "shard owner went away")
    mpiT.Comm_call_errhandler(mpiT.COMM_SELF, code)
    assert calls == [(mpiT.COMM_SELF, code)]
    assert mpiT.Error_class(code) == cls and mpiT.Error_string(code) == "shard owner went away"
    assert mpiT.Error_class(mpiT.ERR_TRUNCATE) == mpiT.ERR_TRUNCATE
    mpiT.Comm_set_errhandler(mpiT.COMM_SELF, mpiT.ERRORS_RETURN)
    mpiT.Comm_call_errhandler(mpiT.COMM_SELF, code)                  # returns
    assert len(calls) == 1
    weh = mpiT.Win_create_errhandler(lambda w, c: calls.append(("win", c)))
    with pytest.raises(mpiT.MPIError):                                # a window handler on a communicator
        mpiT.Comm_set_errhandler(mpiT.COMM_SELF, weh)
    assert mpiT.Errhandler_free(eh) == mpiT.ERRHANDLER_NULL
    mpiT.Comm_set_errhandler(mpiT.COMM_SELF, mpiT.ERRORS_ARE_FATAL)


def test_dims_create_and_group_ranges():
    assert mpiT.Dims_create(12, [0, 0]) == [4, 3]
    assert mpiT.Dims_create(16, [0, 2, 0]) == [4, 2, 2]
    assert mpiT.Dims_create(7, [0, 0, 0]) == [7, 1, 1]
    with pytest.raises(mpiT.MPIError):
        mpiT.Dims_create(10, [3, 0])
    g = mpiT.Group(list(range(100, 112)))
    assert mpiT.Group_range_incl(g, [(0, 6, 3), (11, 9, -1)]).ranks == [100, 103, 106, 111, 110, 109]
    assert mpiT.Group_range_excl(g, [(0, 11, 2)]).ranks == [101, 103, 105, 107, 109, 111]


def test_reduce_local_buffer_attach_op_commutative():
    a, b = torch.tensor([1., 5., 3.]), torch.tensor([4., 2., 6.])
    mpiT.Reduce_local(a, b, 3, mpiT.FLOAT, mpiT.MAX)
    assert b.tolist() == [4., 5., 6.]
    i, j = torch.tensor([6, 3], dtype=torch.int32), torch.tensor([3, 5], dtype=torch.int32)
    mpiT.Reduce_local(i, j, 2, mpiT.INT, mpiT.BAND)
    assert j.tolist() == [2, 1]
    op = mpiT.Op_create(lambda x, y: y.sub_(x), commute=False)
    assert not mpiT.Op_commutative(op) and mpiT.Op_commutative(mpiT.SUM)
    buf = mpiT.Alloc_mem(1 << 12)
    mpiT.Buffer_attach(buf)
    with pytest.raises(mpiT.MPIError):
        mpiT.Buffer_attach(buf)
    got, n = mpiT.Buffer_detach()
    assert got is buf and n == 1 << 12
    mpiT.Pcontrol(2)
    assert os.environ["MPIT_PROFILE_LEVEL"] == "2"
    mpiT.Pcontrol(1)


def test_generalized_request_completes_from_another_thread():
    seen = {}
    req = mpiT.Grequest_start(lambda state, st: seen.setdefault("query", state),
                              lambda state: seen.setdefault("free", state), None, "ctx")
    assert not mpiT.Test(req) and not mpiT.Request_get_status(req)
    threading.Timer(0.05, mpiT.Grequest_complete, args=(req,)).start()
    mpiT.Wait(req)
    assert seen == {"query": "ctx", "free": "ctx"} and mpiT.Test(req)


# ---- MPI-IO on a COMM_SELF file ------------------------------------------------------------------------------------
def test_file_views_pointers_and_sizes(tmp_path):
    path = str(tmp_path / "params.bin")
    fh = mpiT.File_open(mpiT.COMM_SELF, path, mpiT.MODE_CREATE | mpiT.MODE_RDWR)
    data = torch.arange(64, dtype=torch.float32)
    st = mpiT.Status()
    assert mpiT.File_write(fh, data, 64, mpiT.FLOAT, st) == 256 and st.count_bytes == 256
    assert mpiT.File_get_size(fh) == 256 and mpiT.File_get_position(fh) == 256      # default etype is BYTE
    # a view that sees every other pair of floats after a 16-byte header
    pair_of_four = mpiT.Type_commit(mpiT.Type_create_resized(mpiT.Type_contiguous(2, mpiT.FLOAT), 0, 16))
    mpiT.File_set_view(fh, 16, mpiT.FLOAT, pair_of_four)
    assert mpiT.File_get_view(fh) == (16, mpiT.FLOAT, pair_of_four, "native") and mpiT.File_get_position(fh) == 0
    got = torch.zeros(6)
    mpiT.File_read(fh, got, 6, mpiT.FLOAT)
    assert got.tolist() == [4, 5, 8, 9, 12, 13]
    assert mpiT.File_get_position(fh) == 6 and mpiT.File_get_byte_offset(fh, 6) == 16 + 3 * 16
    mpiT.File_seek(fh, -2, mpiT.SEEK_CUR)
    mpiT.File_read(fh, got, 2, mpiT.FLOAT)
    assert got[:2].tolist() == [12, 13]
    mpiT.File_seek(fh, 0, mpiT.SEEK_END)                       # (256-16)/16 tiles * 2 floats
    assert mpiT.File_get_position(fh) == 30
    # write through the view with a strided memory datatype: every third element of src
    src = torch.arange(100, 130, dtype=torch.float32)
    third = mpiT.Type_commit(mpiT.Type_vector(4, 1, 3, mpiT.FLOAT))
    mpiT.File_write_at(fh, 2, src, 1, third)
    whole = torch.zeros(64)
    mpiT.File_set_view(fh, 0, mpiT.FLOAT, mpiT.FLOAT)
    assert mpiT.File_read_at_all(fh, 0, whole, 64, mpiT.FLOAT) == 256
    want = data.clone()
    want[[8, 9, 12, 13]] = torch.tensor([100., 103., 106., 109.])
    assert torch.equal(whole, want)
    # reading past the end is short, not an error
    tail = torch.full((8,), -1.0)
    assert mpiT.File_read_at(fh, 60, tail, 8, mpiT.FLOAT, st) == 16 and mpiT.Get_count(st, mpiT.FLOAT) == 4
    assert tail.tolist() == [60, 61, 62, 63, -1, -1, -1, -1]
    mpiT.File_set_size(fh, 128)
    assert mpiT.File_get_size(fh) == 128
    mpiT.File_preallocate(fh, 512)
    assert mpiT.File_get_size(fh) == 512
    assert mpiT.File_get_amode(fh) & mpiT.MODE_RDWR and mpiT.File_get_type_extent(fh, pair_of_four) == 16
    mpiT.File_set_atomicity(fh, True)
    assert mpiT.File_get_atomicity(fh)
    mpiT.File_write_at(fh, 0, torch.ones(2), 2, mpiT.FLOAT)    # under the record lock
    mpiT.File_sync(fh)
    info = mpiT.Info_create()
    mpiT.Info_set(info, "cb_nodes", "1")
    mpiT.File_set_info(fh, info)
    assert mpiT.Info_get(mpiT.File_get_info(fh), "cb_nodes") == ("1", True)
    assert mpiT.File_close(fh) is mpiT.FILE_NULL
    assert not os.path.exists(str(tmp_path / ".params.bin.shfp"))
    with pytest.raises(mpiT.MPIError) as e:
        mpiT.File_open(mpiT.COMM_SELF, path, mpiT.MODE_CREATE | mpiT.MODE_EXCL | mpiT.MODE_WRONLY)
    assert e.value.code == mpiT.ERR_FILE_EXISTS
    mpiT.File_delete(path)
    with pytest.raises(mpiT.MPIError) as e:
        mpiT.File_open(mpiT.COMM_SELF, path, mpiT.MODE_RDONLY)
    assert e.value.code == mpiT.ERR_NO_SUCH_FILE
    with pytest.raises(mpiT.MPIError) as e:
        mpiT.File_open(mpiT.COMM_SELF, path, mpiT.MODE_RDONLY | mpiT.MODE_CREATE)
    assert e.value.code == mpiT.ERR_AMODE


def test_file_shared_pointer_nonblocking_split_and_datareps(tmp_path):
    path = str(tmp_path / "log.bin")
    fh = mpiT.File_open(mpiT.COMM_SELF, path, mpiT.MODE_CREATE | mpiT.MODE_RDWR | mpiT.MODE_DELETE_ON_CLOSE)
    mpiT.File_set_view(fh, 0, mpiT.INT, mpiT.INT)
    a = torch.arange(4, dtype=torch.int32)
    mpiT.File_write_shared(fh, a, 4, mpiT.INT)
    mpiT.File_write_ordered(fh, a + 10, 4, mpiT.INT)
    assert mpiT.File_get_position_shared(fh) == 8
    r1 = mpiT.File_iwrite_shared(fh, a + 20, 4, mpiT.INT)
    r2 = mpiT.File_iwrite_at(fh, 12, a + 30, 4, mpiT.INT)
    st = mpiT.Status()
    mpiT.Wait(r1, st)
    assert st.count_bytes == 16
    while not mpiT.Test(r2):
        pass
    mpiT.File_seek_shared(fh, 0, mpiT.SEEK_SET)
    out = torch.zeros(16, dtype=torch.int32)
    mpiT.File_read_ordered_begin(fh, out, 16, mpiT.INT)
    with pytest.raises(mpiT.MPIError):                                      # only one split collective at a time
        mpiT.File_read_all_begin(fh, out, 1, mpiT.INT)
    assert mpiT.File_read_ordered_end(fh, out, st) == 64
    assert out.tolist() == list(range(4)) + list(range(10, 14)) + list(range(20, 24)) + list(range(30, 34))
    half = torch.zeros(8, dtype=torch.int32)
    mpiT.File_seek(fh, 8, mpiT.SEEK_SET)
    r3 = mpiT.File_iread(fh, half, 8, mpiT.INT)
    assert mpiT.File_get_position(fh) == 16                                 # advanced at the call
    mpiT.Wait(r3)
    assert half.tolist() == out[8:].tolist()
    # external32 stores big-endian; the raw bytes prove it
    mpiT.File_set_view(fh, 64, mpiT.INT, mpiT.INT, "external32")
    mpiT.File_write_at(fh, 0, torch.tensor([1, 258], dtype=torch.int32), 2, mpiT.INT)
    back = torch.zeros(2, dtype=torch.int32)
    mpiT.File_read_at(fh, 0, back, 2, mpiT.INT)
    assert back.tolist() == [1, 258]
    mpiT.File_set_view(fh, 64, mpiT.BYTE, mpiT.BYTE)
    raw = torch.zeros(8, dtype=torch.uint8)
    mpiT.File_read_at(fh, 0, raw, 8, mpiT.BYTE)
    assert raw.tolist() == [0, 0, 0, 1, 0, 0, 1, 2]
    # a user representation: bytes stored inverted
    mpiT.Register_datarep("inverted", lambda s, d, c, x: 255 - s, lambda s, d, c, x: 255 - s, None, None)
    with pytest.raises(mpiT.MPIError):
        mpiT.Register_datarep("inverted", None, None)
    mpiT.File_set_view(fh, 128, mpiT.BYTE, mpiT.BYTE, "inverted")
    mpiT.File_write_at(fh, 0, torch.tensor([1, 2, 3], dtype=torch.uint8), 3, mpiT.BYTE)
    chk = torch.zeros(3, dtype=torch.uint8)
    mpiT.File_read_at(fh, 0, chk, 3, mpiT.BYTE)
    assert chk.tolist() == [1, 2, 3]
    mpiT.File_set_view(fh, 128, mpiT.BYTE, mpiT.BYTE)
    mpiT.File_read_at(fh, 0, chk, 3, mpiT.BYTE)
    assert chk.tolist() == [254, 253, 252]
    # sequential mode forbids explicit offsets; read-only forbids writes
    mpiT.File_close(fh)
    assert not os.path.exists(path)
    open(path, "wb").write(bytes(range(16)))
    ro = mpiT.File_open(mpiT.COMM_SELF, path, mpiT.MODE_RDONLY | mpiT.MODE_SEQUENTIAL)
    with pytest.raises(mpiT.MPIError) as e:
        mpiT.File_read_at(ro, 0, chk, 3, mpiT.BYTE)
    assert e.value.code == mpiT.ERR_UNSUPPORTED_OPERATION
    with pytest.raises(mpiT.MPIError) as e:
        mpiT.File_write_shared(ro, chk, 3, mpiT.BYTE)
    assert e.value.code == mpiT.ERR_READ_ONLY
    mpiT.File_read_shared(ro, chk, 3, mpiT.BYTE)
    mpiT.File_read_shared(ro, chk, 3, mpiT.BYTE)
    assert chk.tolist() == [3, 4, 5]
    mpiT.File_close(ro)


REF_SPECS = "/root/reference/specs"


@pytest.mark.skipif(not os.path.isdir(REF_SPECS), reason="reference specs absent")
def test_every_function_in_the_reference_specs_is_provided():
    import importlib.util
    here = os.path.dirname(os.path.abspath(__file__))
    spec = importlib.util.spec_from_file_location("mpi_coverage", os.path.join(here, "..", "tools", "mpi_coverage.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    rows = mod.coverage(REF_SPECS)
    missing = [n for v in rows.values() for n, have in v if not have]
    assert sum(len(v) for v in rows.values()) == 270 and not missing, missing
