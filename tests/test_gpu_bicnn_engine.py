"""GPU numerics: BiCNN kernel engine (tcgen05 GEMMs + fused epilogues + bicnn.cu) vs the torch model on CPU."""
import pytest
import torch

from mpit_b200.models.bicnn import BiCNN, gesd, margin_ranking, pad_sentences

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-12))


@pytest.mark.parametrize("dims", [(100, 200, 3000), (32, 64, 128)])
def test_bicnn_engine_forward_backward_matches_torch(dims):
    from mpit_b200.models.bicnn_engine import BiCnnEngine
    E, H, F = dims
    torch.manual_seed(0)
    V, S = 300, 12
    model = BiCNN(V, E, H, F, 2)
    sents = [torch.randint(0, V, (int(n),)) for n in torch.randint(4, 30, (S,))]
    tokens, lens = pad_sentences(sents)
    # triples: (q, a+, a-) = sentences (0,1,2), (3,4,5), (6,7,8), (9,10,11)
    qi = torch.tensor([0, 0, 3, 3, 6, 6, 9, 9], dtype=torch.int32)
    ai = torch.tensor([1, 2, 4, 5, 7, 8, 10, 11], dtype=torch.int32)
    margin = 0.5      # large margin: every triple is active and contributes gradient
    # ---- oracle
    y_ref = model.encode(tokens, lens)
    s_ref = gesd(y_ref[qi.long()], y_ref[ai.long()])
    loss_ref = margin_ranking(s_ref[0::2], s_ref[1::2], margin).sum()
    loss_ref.backward()
    g_ref = torch.cat([p.grad.reshape(-1) for p in model.param_list()])
    theta = torch.cat([p.detach().reshape(-1) for p in model.param_list()])
    # ---- engine
    eng = BiCnnEngine(V, S, 32, E, H, F, 2, margin=margin)
    th = theta.cuda()
    eng.load_flat(th)
    eng.set_sentences(tokens.cuda().int(), lens.cuda().int())
    y = eng.encode()
    torch.cuda.synchronize()
    eng.check()
    assert _rel(y.cpu(), y_ref.detach()) < 3e-3, _rel(y.cpu(), y_ref.detach())
    sim = eng.gesd(qi.cuda(), ai.cuda())
    assert torch.allclose(sim.cpu(), s_ref.detach(), atol=2e-3)
    grad = torch.zeros_like(th)
    loss = eng.loss_backward(qi.cuda(), ai.cuda(), grad)
    eng.flat_grad(grad)
    torch.cuda.synchronize()
    eng.check()
    assert abs(float(loss) - float(loss_ref)) < 5e-3 * max(1.0, float(loss_ref))
    g = grad.cpu()
    n_lk = V * E
    names = [("lookup", 0, n_lk), ("lin.w", n_lk, n_lk + H * E), ("lin.b", n_lk + H * E, n_lk + H * E + H),
             ("conv.w", n_lk + H * E + H, n_lk + H * E + H + F * 2 * H), ("conv.b", n_lk + H * E + H + F * 2 * H, g.numel())]
    for name, lo, hi in names:
        # TF32 forward can flip an arg-max among near-equal frames; a loose bound per block, tight overall
        assert _rel(g[lo:hi], g_ref[lo:hi]) < 0.08, (name, _rel(g[lo:hi], g_ref[lo:hi]))
    assert _rel(g, g_ref) < 0.05, _rel(g, g_ref)


def test_kernel_trainer_feval_matches_torch_trainer_and_learns():
    """The B200 feval selects the same negatives and produces the same loss / gradient as the torch
    restatement, and a few epochs of single-worker Adam raise top-1 accuracy."""
    from mpit_b200.launch.plaunch import parse
    from mpit_b200.optim import reference as R
    from mpit_b200.workloads import bicnn as W
    corpus = W.synthetic_corpus(vocab=300, n_answers=40, n_train=96, n_eval=40, pool=8, emb_dim=32, seed=1)
    opt = parse(["-embeddingDim", "32", "-wordHiddenDim", "64", "-numFilters", "128", "-maxnegsample", "6",
                 "-batchSize", "4", "-L2reg", "0.0001", "-gradClip", "0.5", "-margin", "0.05"])
    torch.manual_seed(3)
    cpu = W.BiCNNTrainer(corpus, opt, "cpu", rank=0)
    gpu = W.BiCNNKernelTrainer(corpus, opt, "cuda", rank=0)
    gpu.parameters.copy_(cpu.parameters)
    cpu.batch = gpu.batch = [0, 1, 2, 3]
    f_c, g_c = cpu.feval(cpu.parameters)
    f_g, g_g = gpu.feval(gpu.parameters)
    assert abs(f_c - f_g) < 5e-3 * max(1.0, abs(f_c)), (f_c, f_g)
    assert _rel(g_g.cpu(), g_c) < 0.08, _rel(g_g.cpu(), g_c)
    before = gpu.test(corpus.valid, "valid")
    cfg = {"lr": 0.01, "beta1": 0.9, "beta2": 0.999, "epsilon": 1e-8}
    for epoch in range(4):
        order = torch.randperm(len(corpus.train)).tolist()
        for t in range(0, len(order), 4):
            gpu.batch = order[t:t + 4]
            R.adamsingle(gpu.feval, gpu.parameters, cfg)
    after = gpu.test(corpus.valid, "valid")
    gpu.eng.check()
    assert after > before + 0.1, (before, after)
