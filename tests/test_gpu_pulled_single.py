"""The pull-fused GEMM on one GPU (one owner = this GPU): both rings of the kernel, every M-tiles-per-CTA variant, the
data gradient with and without split-K, the weight gradient and the push -- against plain PyTorch fp32.  The
multi-GPU form of the same checks is ``case_pulled_linear`` in tests/mp_gpu_cases.py."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a - b).norm() / b.norm())


@pytest.fixture(scope="module")
def heap():
    from mpit_b200.parallel.symm import SymmetricHeap
    return SymmetricHeap(torch.device("cuda", 0))


@pytest.mark.parametrize("M,K,N", [(200, 256, 512), (512, 1024, 1024), (700, 64, 128), (128, 2048, 256)])
def test_forward_every_tiling_matches_torch(heap, M, K, N):
    from mpit_b200.ops.pulled_linear import PulledLinear
    dev = torch.device("cuda", 0)
    gen = torch.Generator().manual_seed(M + K)
    W, b = torch.randn(N, K, generator=gen) * 0.05, torch.randn(N, generator=gen) * 0.1
    x = torch.randn(M, K, generator=gen)
    lin = PulledLinear(heap, K, N, M, act=1)
    lin.load_rows(W.to(dev))
    ref = torch.relu(x @ W.t() + b)
    for mt in (0, 1, 2, 4):
        y = lin.forward(x.to(dev), b.to(dev), mt=mt)
        torch.cuda.synchronize()
        lin.check()
        assert _rel(y.cpu(), ref) < 3e-3, (mt,)
    y = lin.forward(x.to(dev), b.to(dev), fused=False)
    torch.cuda.synchronize()
    assert _rel(y.cpu(), ref) < 3e-3


@pytest.mark.parametrize("M,K,N", [(200, 256, 512), (512, 1024, 4096), (700, 64, 256)])
def test_data_gradient_with_split_k_matches_torch(heap, M, K, N):
    from mpit_b200.ops.pulled_linear import PulledLinear
    dev = torch.device("cuda", 0)
    gen = torch.Generator().manual_seed(M + N)
    W = torch.randn(N, K, generator=gen) * 0.05
    x = torch.randn(M, K, generator=gen)
    dy = torch.randn(M, N, generator=gen) * 0.1
    lin = PulledLinear(heap, K, N, M, act=0)
    lin.load_rows(W.to(dev))
    lin.forward(x.to(dev))
    ref = dy @ W
    nkb = N // 32
    for mt, ks in ((0, 0), (1, 1), (2, 2), (4, 4), (1, 8), (4, nkb)):
        if nkb % max(ks, 1):
            continue
        lin.dx.fill_(float("nan"))                              # split-K must zero what it adds into
        dx = lin.backward(dy.to(dev), mt=mt, ksplit=ks)
        torch.cuda.synchronize()
        lin.check()
        assert _rel(dx.cpu(), ref) < 3e-3, (mt, ks)
    gw = lin.weight_grad()
    torch.cuda.synchronize()
    assert _rel(gw.cpu(), dy.t() @ x) < 3e-3
    lin.push(-0.5)
    torch.cuda.synchronize()
    want = W - 0.5 * (dy.t() @ x)
    assert _rel(lin.weight_rows().cpu(), want) < 3e-3
    assert torch.allclose(lin.weight_rows_t().cpu(), lin.weight_rows().cpu().t(), atol=1e-5)


def test_split_k_arguments_are_validated(heap):
    from mpit_b200._native import kernels, ptr, cur_stream
    from mpit_b200.ops.pulled_linear import PulledLinear
    lin = PulledLinear(heap, 256, 512, 128, act=0)
    k = kernels()
    bias = torch.zeros(512, device="cuda")
    _, ta = lin._operand(torch.zeros(128, 256, device="cuda"), 256, lin.x)
    _, ta_dy = lin._operand(torch.zeros(128, 512, device="cuda"), 512, lin.dy)
    # split-K only for the data gradient, without bias / activation, and only with a k-block count it divides
    assert k.mpit_umma_pulled_gemm(ta, lin._maps, 1, 512, 0, 128, 256, 512, 0, None, ptr(lin.y), ptr(lin.err),
                                   lin.block_n, 1, 2, cur_stream()) != 0
    assert k.mpit_umma_pulled_gemm(ta_dy, lin._maps_t, 1, 512, 1, 128, 512, 256, 0, ptr(bias), ptr(lin.dx),
                                   ptr(lin.err), lin.block_k, 1, 2, cur_stream()) != 0
    assert k.mpit_umma_pulled_gemm(ta_dy, lin._maps_t, 1, 512, 1, 128, 512, 256, 0, None, ptr(lin.dx),
                                   ptr(lin.err), lin.block_k, 1, 3, cur_stream()) != 0


def test_inputs_that_cannot_be_read_in_place_are_staged(heap):
    from mpit_b200.ops.pulled_linear import PulledLinear
    dev = torch.device("cuda", 0)
    gen = torch.Generator().manual_seed(5)
    M, K, N = 130, 64, 128
    W = torch.randn(N, K, generator=gen) * 0.05
    lin = PulledLinear(heap, K, N, M, act=0)
    lin.load_rows(W.to(dev))
    xt = torch.randn(K, M, generator=gen).to(dev).t()            # a transposed view: not contiguous
    assert not xt.is_contiguous()
    y = lin.forward(xt)
    torch.cuda.synchronize()
    assert _rel(y.cpu(), xt.cpu() @ W.t()) < 3e-3
    dy = torch.randn(M, N, generator=gen).double().to(dev)       # wrong dtype
    dx = lin.backward(dy)
    gw = lin.weight_grad()
    torch.cuda.synchronize()
    lin.check()
    assert _rel(dx.cpu(), dy.float().cpu() @ W) < 3e-3
    assert _rel(gw.cpu(), dy.float().cpu().t() @ xt.cpu()) < 3e-3
