"""The allocation-free weight-gradient entry point, called through the C ABI.

``mpit_umma_wgrad_noalloc`` exists for users of the library who cannot hand it a workspace: a C
program or other-language binding that has no device allocator, a process whose GPU memory is
already entirely owned by a framework's caching allocator, or a loop that cannot afford the
device-wide synchronisation of ``cudaMalloc``/``cudaFree``.  Its split-K scratch is a fixed 8 MiB
array inside ``libmpit_kernels.so``.  This test drives it with raw ``ctypes`` calls and checks that
it reproduces the regular entry points (``mpit_umma_wgrad`` + ``mpit_wgrad_reduce`` with a
caller-owned workspace), twice in a row on different layers so that the second call runs on top of
whatever the first one left in the shared scratch.  It lives in a file that sorts last so that a
``pytest -x`` run has exercised everything else before it gets here.
"""
import ctypes

import pytest
import torch

from mpit_b200._native import kernels, ptr, cur_stream
from mpit_b200.models.cifar_cnn import CifarEngine, TorchCifarCNN, flat_from_torch, random_crop_offsets, NUM_CLASSES

pytestmark = pytest.mark.gpu

MPIT_E_SCRATCH = -5
SCRATCH_FLOATS = 2 * 1024 * 1024   # MPIT_NOALLOC_SCRATCH_FLOATS in include/mpit_b200.h
TOL = 3e-3                         # the layer-wise tolerance of tests/test_gpu_cifar_engine.py


def _rel_l2(a, b):
    return float((a - b).norm() / (b.norm() + 1e-12))


def _engine(B=32):
    torch.manual_seed(11)
    eng = CifarEngine(B, impl="umma", seed=5)
    eng.load_flat(flat_from_torch(TorchCifarCNN()).cuda())
    eng.set_batch(torch.randint(0, 256, (B, 3, 32, 32), dtype=torch.uint8),
                  torch.randint(0, NUM_CLASSES, (B,), dtype=torch.int32), random_crop_offsets(B))
    eng.fwd_bwd()          # fills activations, dz and -- through the regular entry points -- gpk
    torch.cuda.synchronize()
    eng._umma.check()
    return eng


def _call_noalloc(eng, layer, out):
    """Raw C-ABI call: int mpit_umma_wgrad_noalloc(tmap_x, tmap_dz, geom11, gw, ldw, crop, err,
    block_n, nsplit, stream)."""
    l = eng._umma.wgrad.launches[layer]
    return kernels().mpit_umma_wgrad_noalloc(l.tmap_x, l.tmap_dz, l.geom, ctypes.c_void_p(out.data_ptr()),
                                             ctypes.c_int(l.K), ptr(l.crop), ptr(eng._umma.err),
                                             ctypes.c_int(l.block_n), ctypes.c_int(l.nsplit), cur_stream())


def test_noalloc_wgrad_matches_regular_entry_point_twice():
    eng = _engine()
    k = kernels()
    # first call: conv2 (the largest scratch user), second call: conv1 (different shape, crop path),
    # third: conv2 again with a changed dz -- each on top of the previous call's leftovers
    for it, layer in enumerate((1, 0, 1)):
        L = eng.layers[layer]
        if layer == 1:
            eng.dz[1].mul_(1.5)
        l = eng._umma.wgrad.launches[layer]
        assert l.nsplit * l.m_pad * l.N <= SCRATCH_FLOATS
        # regular path with a caller-owned workspace
        part = torch.empty(l.nsplit * l.m_pad * l.N, dtype=torch.float32, device="cuda")
        want = torch.full((L.wsize,), float("nan"), device="cuda")
        assert k.mpit_umma_wgrad(l.tmap_x, l.tmap_dz, l.geom, ptr(part), ptr(l.crop), ptr(eng._umma.err),
                                 l.block_n, l.nsplit, cur_stream()) == 0
        assert k.mpit_wgrad_reduce(ptr(part), ptr(want), l.nsplit, l.m_pad, l.K, l.N, l.K, cur_stream()) == 0
        # allocation-free path
        got = torch.full((L.wsize,), float("nan"), device="cuda")
        torch.cuda.synchronize()
        before, free_before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
        assert _call_noalloc(eng, layer, got) == 0
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        if it > 0:   # after the first call's one-time set-up the device's free memory does not move
            assert torch.cuda.mem_get_info()[0] == free_before
        eng._umma.check()
        assert torch.isfinite(got).all() and torch.isfinite(want).all()
        assert float(want.abs().max()) > 0
        assert _rel_l2(got, want) < TOL, (layer, _rel_l2(got, want))


def test_noalloc_wgrad_rejects_oversized_request(capfd):
    eng = _engine(B=8)
    l = eng._umma.wgrad.launches[1]
    out = torch.zeros(eng.layers[1].wsize, device="cuda")
    too_many = SCRATCH_FLOATS // (l.m_pad * l.N) + 1
    rc = kernels().mpit_umma_wgrad_noalloc(l.tmap_x, l.tmap_dz, l.geom, ctypes.c_void_p(out.data_ptr()),
                                           ctypes.c_int(l.K), ptr(l.crop), ptr(eng._umma.err),
                                           ctypes.c_int(l.block_n), ctypes.c_int(too_many), cur_stream())
    assert rc == MPIT_E_SCRATCH
    assert "scratch" in capfd.readouterr().err
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0   # nothing was launched
