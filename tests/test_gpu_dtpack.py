"""GPU: derived-datatype gather / scatter on the device (csrc/kernels/dtpack.cu) against the host engine, which the
CPU tests check against NumPy slicing."""
import numpy as np
import pytest
import torch

from mpit_b200 import msg as mpiT

pytestmark = pytest.mark.gpu


def _types():
    col2 = mpiT.Type_commit(mpiT.Type_create_resized(mpiT.Type_vector(64, 2, 96, mpiT.FLOAT), 0, 8))
    sub = mpiT.Type_commit(mpiT.Type_create_subarray([32, 48, 16], [8, 12, 8], [3, 5, 4], mpiT.ORDER_C, mpiT.FLOAT))
    idx = mpiT.Type_commit(mpiT.Type_indexed([3, 1, 7], [40, 0, 9], mpiT.DOUBLE))
    odd = mpiT.Type_commit(mpiT.Type_create_struct([1, 3, 1], [0, 5, 12], [mpiT.INT, mpiT.CHAR, mpiT.SHORT]))
    big = mpiT.Type_commit(mpiT.Type_vector(1024, 256, 1024, mpiT.FLOAT))          # 1 MiB gathered from 4 MiB
    return {"col2": (col2, 5, 64 * 96), "sub": (sub, 2, 2 * 32 * 48 * 16), "idx": (idx, 4, 4 * 43 * 2),
            "odd": (odd, 9, 9 * 4), "big": (big, 1, 1024 * 1024)}


@pytest.mark.parametrize("name", ["col2", "sub", "idx", "odd", "big"])
def test_device_pack_and_unpack_match_the_host_engine(name):
    dtype, count, nfloats = _types()[name]
    g = torch.Generator().manual_seed(3)
    host = torch.randint(0, 2 ** 31 - 1, (nfloats,), generator=g, dtype=torch.int32).view(torch.float32)
    dev = host.cuda()
    want = np.zeros(mpiT.Pack_size(count, dtype), dtype=np.uint8)
    mpiT.Pack(host, count, dtype, want, 0)
    got = np.zeros_like(want)
    mpiT.Pack(dev, count, dtype, got, 0)
    assert np.array_equal(got, want)
    # scatter into a different buffer: untouched bytes stay, touched bytes equal the host result
    base = torch.full((nfloats,), -1.0)
    h2, d2 = base.clone(), base.clone().cuda()
    mpiT.Unpack(want, 0, h2, count, dtype)
    mpiT.Unpack(want, 0, d2, count, dtype)
    torch.cuda.synchronize()
    assert torch.equal(d2.cpu().view(torch.int32), h2.view(torch.int32))


def test_misaligned_device_views_fall_back_to_narrow_words():
    col = mpiT.Type_commit(mpiT.Type_vector(16, 4, 10, mpiT.FLOAT))
    host = torch.arange(200, dtype=torch.float32)
    dev = host.cuda()
    for off in (0, 1, 2, 3):                                   # 4-byte steps off a 16-byte boundary
        want = np.zeros(mpiT.Pack_size(1, col), dtype=np.uint8)
        got = np.zeros_like(want)
        mpiT.Pack(host[off:], 1, col, want, 0)
        mpiT.Pack(dev[off:], 1, col, got, 0)
        assert np.array_equal(got, want), off


def test_file_write_and_read_of_a_device_tensor_through_a_view(tmp_path):
    """A CUDA tensor written through a strided memory datatype and read back into another CUDA tensor."""
    path = str(tmp_path / "dev.bin")
    fh = mpiT.File_open(mpiT.COMM_SELF, path, mpiT.MODE_CREATE | mpiT.MODE_RDWR)
    x = torch.arange(64 * 64, dtype=torch.float32).view(64, 64).cuda()
    tile = mpiT.Type_commit(mpiT.Type_create_subarray([64, 64], [16, 32], [8, 16], mpiT.ORDER_C, mpiT.FLOAT))
    assert mpiT.File_write_at(fh, 0, x, 1, tile) == 16 * 32 * 4
    flat = torch.zeros(16, 32)
    mpiT.File_read_at(fh, 0, flat, 16 * 32 * 4, mpiT.BYTE)
    assert torch.equal(flat, x[8:24, 16:48].cpu())
    y = torch.zeros(64, 64, device="cuda")
    mpiT.File_read_at(fh, 0, y, 1, tile)
    torch.cuda.synchronize()
    want = torch.zeros(64, 64)
    want[8:24, 16:48] = x[8:24, 16:48].cpu()
    assert torch.equal(y.cpu(), want)
    mpiT.File_close(fh)
