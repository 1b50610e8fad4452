"""GPU tests of the device-resident BiCNN training step (workloads.bicnn.BiCNNDeviceTrainer): batch assembly with
negative sampling and first-violator selection on the device against a host recomputation
(/root/reference/BiCNN/bicnn.lua:321-359), the gradient against the host-driven kernel path, and a short training run
through the asynchronous parameter server with server-side Adam (BiCNN/optim-adam.lua + pserver.lua:140-155)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

ARGS = ["-embeddingDim", "32", "-wordHiddenDim", "64", "-numFilters", "128", "-maxnegsample", "6", "-batchSize", "4",
        "-L2reg", "0.0001", "-gradClip", "0.5", "-margin", "0.05", "-optimization", "adam", "-lrAdam", "0.01",
        "-stepDivAdam", "0"]


def _rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-12))


def _make(extra=(), graphs=False, server="inline"):
    from mpit_b200.launch.plaunch import parse
    from mpit_b200.parallel.symm import SymmetricHeap
    from mpit_b200.workloads import bicnn as W
    corpus = W.synthetic_corpus(vocab=300, n_answers=40, n_train=96, n_eval=40, pool=8, emb_dim=32, seed=1)
    opt = parse(ARGS + list(extra))
    torch.manual_seed(3)
    heap = SymmetricHeap(torch.device("cuda", 0))
    tr = W.BiCNNDeviceTrainer(corpus, opt, "cuda:0", 0, heap, server=server, use_graphs=graphs)
    return W, corpus, opt, tr


def test_device_batch_assembly_selection_and_gradient():
    W, corpus, opt, tr = _make()
    eng = tr.eng
    B, K = opt["batchSize"], opt["maxnegsample"]
    batch = [5, 17, 2]                                  # one empty slot
    tr._stage_batch(batch)
    eng.train_fwd_bwd()
    torch.cuda.synchronize()
    eng.check()
    tok, lens = eng.tokens.cpu(), eng.lens.cpu()
    negs = eng.neg_idx.cpu().view(B, K)
    for b, idx in enumerate(batch):
        labels, q, a = corpus.train[idx]
        assert tok[b, :q.numel()].tolist() == q.tolist() and int(lens[b]) == q.numel()
        assert tok[B + b, :a.numel()].tolist() == a.tolist() and int(lens[B + b]) == a.numel()
        for k in range(K):
            j = int(negs[b, k])
            assert 0 <= j < len(corpus.answers) and j not in labels              # bicnn.lua:327-334
            ans = corpus.answers[j]
            s = 2 * B + b * K + k
            assert tok[s, :ans.numel()].tolist() == ans.tolist() and int(lens[s]) == ans.numel()
    assert len({tuple(r) for r in negs[:3].tolist()}) > 1, "every sample draws its own candidates"
    assert int(eng.frames[3]) == 0 and int(eng.frames[B + 3]) == 0              # the empty slot encodes nothing
    # first violator, recomputed from the similarities the device produced
    sim = eng.sim_all.cpu()
    ai2 = eng.ai2.cpu()
    sim2 = eng.sim2.cpu()
    n_active = 0
    for b in range(B):
        first = -1
        if b < len(batch):
            for k in range(K):
                if float(sim[b]) - float(sim[B + b * K + k]) < opt["margin"]:
                    first = k
                    break
        if first >= 0:
            n_active += 1
            assert int(ai2[2 * b]) == B + b and int(ai2[2 * b + 1]) == 2 * B + b * K + first
            assert float(sim2[2 * b]) == float(sim[b]) and float(sim2[2 * b + 1]) == float(sim[B + b * K + first])
        else:
            assert float(sim2[2 * b]) == 1.0 and float(sim2[2 * b + 1]) == 0.0
    assert n_active >= 1 and int(eng.active) == 1
    assert int(eng.counters[0]) == n_active and int(eng.step_ctr) == 1
    # the same triples through the host-driven path of the engine give the same loss and gradient
    act = [b for b in range(B) if float(sim2[2 * b]) != 1.0 or float(sim2[2 * b + 1]) != 0.0]
    qi = torch.tensor([b for b in act for _ in (0, 1)], dtype=torch.int32, device="cuda")
    ai = torch.stack([ai2[2 * torch.tensor(act)], ai2[2 * torch.tensor(act) + 1]], 1).view(-1).int().cuda()
    g_dev = eng.gflat.clone()
    eng.flat_grad(g_dev)
    loss_dev = float(eng.loss_acc)
    ref = torch.zeros_like(g_dev)
    loss_ref = float(eng.loss_backward(qi, ai, ref))
    eng.flat_grad(ref)
    torch.cuda.synchronize()
    assert abs(loss_dev - loss_ref) < 1e-5 * max(1.0, loss_ref), (loss_dev, loss_ref)
    assert _rel(g_dev, ref) < 1e-5, _rel(g_dev, ref)
    # a second step draws different candidates (the step counter lives on the device)
    eng.train_fwd_bwd()
    torch.cuda.synchronize()
    assert not torch.equal(eng.neg_idx.cpu().view(B, K)[:3], negs[:3])
    tr.close()


def test_every_step_in_flight_keeps_its_own_batch():
    """The host runs ahead of the GPU (the step is a graph launch): the sample indices of step t must still be the ones
    the device sees at step t.  The GPU is held busy while 48 steps' indices are staged; a device-side log records
    what each step would have read."""
    W, corpus, opt, tr = _make([])
    B = opt["batchSize"]
    log = torch.full((48, B), -7, dtype=torch.int32, device="cuda")
    torch.cuda._sleep(int(3e8))                                   # ~0.15 s of GPU time in front of everything
    for i in range(48):
        tr._stage_batch([i, i + 1][:B])
        log[i].copy_(tr.eng.batch)
    torch.cuda.synchronize()
    got = log.cpu()
    for i in range(48):
        assert got[i].tolist()[:2] == [i, i + 1][:B] and all(int(v) == -1 for v in got[i][2:]), (i, got[i].tolist())
    tr.close()


@pytest.mark.parametrize("graphs,server", [(True, "inline"), (False, "inline"), (True, "thread")])
def test_device_training_through_async_server_learns(graphs, server):
    W, corpus, opt, tr = _make(["-epoch", "5"], graphs=graphs, server=server)
    before = tr.test(corpus.valid, "valid")
    out = W.run_device_trainer(tr, False, False)
    after = tr.test(corpus.valid, "valid")
    tr.eng.check()
    assert out["steps"] == 5 * 24
    assert tr.optim.ps.applied == out["steps"] and tr.optim.ps.pushes == out["steps"]
    assert out["loss"][-1] < out["loss"][0], out["loss"]
    assert after > before + 0.1, (before, after)
    st = tr.loss_stats()
    assert st["steps"] == out["steps"] and st["triples"] > 0 and st["l2"] > 0


def test_device_downpour_accumulates_commperiod_steps():
    """Downpour on the same path: -lr * (sum of `commperiod` clamped gradients) lands in the centre every `commperiod`
    steps (optim-downpour.lua:26-45); in between nothing is pushed."""
    W, corpus, opt, tr = _make(["-optimization", "downpour", "-learningRate", "0.05", "-commperiod", "3", "-epoch", "1"])
    out = W.run_device_trainer(tr, False, False)
    assert out["steps"] == 24 and tr.optim.ps.pushes == 8 and tr.optim.ps.applied == 8
    assert torch.isfinite(tr.parameters).all()
