"""BiCNN: model pieces vs direct restatements of the Lua modules, feval semantics, roles, end-to-end run."""
import os

import torch

from mpit_b200.launch.plaunch import assign_roles, parse, OPTIONS
from mpit_b200.launch.run import launch
from mpit_b200.models.bicnn import BiCNN, NormalizeL2, gesd, margin_ranking, pad_sentences
from mpit_b200.workloads import bicnn as W


def test_option_table_matches_reference_count():
    ref_names = [n for n, _, _ in OPTIONS][:58]
    assert len(ref_names) == 58 and "stepDivAdam" in ref_names and "singlemode" in ref_names   # plaunch.lua:10-69
    opt = parse([])
    assert opt["optimization"] == "downpour" and opt["numFilters"] == 3000 and opt["margin"] == 0.02


def test_normalize_backward_equals_full_jacobian():
    """Closed form vs the d x d Jacobian the reference builds (BiCNN/Normalize.lua:53-69)."""
    torch.manual_seed(0)
    x = torch.randn(3, 17, dtype=torch.float64, requires_grad=True)
    g = torch.randn(3, 17, dtype=torch.float64)
    y = NormalizeL2.apply(x)
    y.backward(g)
    normp = (x.detach() ** 2).sum(1) + 1e-10
    norm = normp.sqrt()
    for i in range(3):
        J = (torch.eye(17, dtype=torch.float64) * normp[i] - torch.outer(x.detach()[i], x.detach()[i])) / (normp[i] * norm[i])
        assert torch.allclose(x.grad[i], J @ g[i], atol=1e-12)
    assert torch.allclose(y.detach().norm(dim=1), torch.ones(3, dtype=torch.float64), atol=1e-6)


def test_gesd_and_margin_formulas():
    q, a = torch.tensor([[0.6, 0.8]]), torch.tensor([[1.0, 0.0]])
    dd, l2 = 0.6, ((0.4 ** 2 + 0.8 ** 2) ** 0.5)
    want = 1 / ((1 + l2) * (1 + torch.exp(torch.tensor(-(dd + 1)))))       # bicnn.lua:442
    assert torch.allclose(gesd(q, a), want)
    assert abs(float(margin_ranking(torch.tensor([0.5]), torch.tensor([0.49]), 0.02)) - 0.01) < 1e-6
    assert float(margin_ranking(torch.tensor([0.5]), torch.tensor([0.1]), 0.02)) == 0.0


def test_batched_encoder_equals_per_sentence():
    """Padding + masking must not change a sentence's embedding (the reference runs unpadded sentences)."""
    torch.manual_seed(1)
    m = BiCNN(50, 8, 12, 20, 2)
    sents = [torch.randint(0, 50, (n,)) for n in (3, 7, 5)]
    t, l = pad_sentences(sents)
    batched = m.encode(t, l)
    for i, s in enumerate(sents):
        single = m.encode(s[None, :], torch.tensor([s.numel()]))
        assert torch.allclose(batched[i], single[0], atol=1e-6)
    assert m.conv_w.shape == (20, 24) and sum(p.numel() for p in m.param_list()) == 50 * 8 + 12 * 8 + 12 + 20 * 24 + 20


def test_reference_parameter_count():
    m = BiCNN(22354)      # V from BiCNN/binary_mapWordIdx2WordStr
    assert sum(p.numel() for p in m.param_list()) == 2235400 + 20200 + 1203000


def test_feval_selects_first_violator_and_clamps():
    corpus = W.synthetic_corpus(vocab=120, n_answers=30, n_train=16, n_eval=5, pool=6, emb_dim=8, seed=3)
    opt = parse(["-embeddingDim", "8", "-wordHiddenDim", "10", "-numFilters", "16", "-maxnegsample", "7",
                 "-gradClip", "0.001", "-L2reg", "0.0001"])
    tr = W.BiCNNTrainer(corpus, opt, "cpu", rank=0)
    tr.batch = [0, 1, 2, 3]
    f, g = tr.feval(tr.parameters)
    assert f >= 0 and g.shape == tr.parameters.shape
    assert float(g.abs().max()) <= 0.001 + 1e-9                      # clamp (bicnn.lua:409)
    assert tr.model.lookup.grad.data_ptr() == tr.gradParameters.data_ptr()   # flat aliasing
    # negatives never come from the target set and the draw is reproducible per rank seed
    tr2 = W.BiCNNTrainer(corpus, opt, "cpu", rank=0)
    assert tr2._draw_negatives([5]) == W.BiCNNTrainer(corpus, opt, "cpu", rank=0)._draw_negatives([5])
    assert 5 not in tr2._draw_negatives([5])


def test_training_improves_top1_on_synthetic_corpus():
    from mpit_b200.optim import reference as R
    torch.manual_seed(0)
    corpus = W.synthetic_corpus(vocab=300, n_answers=40, n_train=200, n_eval=60, pool=8, emb_dim=16, seed=1)
    opt = parse(["-embeddingDim", "16", "-wordHiddenDim", "24", "-numFilters", "64", "-maxnegsample", "10",
                 "-batchSize", "8", "-L2reg", "0", "-gradClip", "5"])
    tr = W.BiCNNTrainer(corpus, opt, "cpu", rank=0)
    before = tr.test(corpus.valid, "valid")
    cfg = {"lr": 0.01, "beta1": 0.9, "beta2": 0.999, "epsilon": 1e-8}
    for epoch in range(3):
        order = torch.randperm(len(corpus.train)).tolist()
        for t in range(0, len(order), 8):
            tr.batch = order[t:t + 8]
            R.adamsingle(tr.feval, tr.parameters, cfg)
    after = tr.test(corpus.valid, "valid")
    assert after > before + 0.15, (before, after)


def test_roles_follow_reference_rules():
    opt = parse(["-testerlast"])
    roles = [assign_roles(r, 5, opt)[0] for r in range(5)]
    assert roles == ["pt", "ps", "pt", "ps", "pe"]                    # (i+1) % masterFreq != 0 -> trainer
    _, sranks, cranks, tranks = assign_roles(0, 5, opt)
    assert sranks == [1, 3] and cranks == [0, 2, 4] and tranks == {4}
    opt = parse(["-testerfirst"])
    assert [assign_roles(r, 5, opt)[0] for r in range(5)] == ["pe", "pt", "ps", "pt", "ps"]


def test_plaunch_parallel_adam_with_tester(tmp_path):
    """BASELINE.json config 4 in miniature: server-side Adam, one trainer, one tester, Torch7 checkpoints."""
    prefix = str(tmp_path / "bicnn")
    rc = launch(3, ["mpit_b200.launch.plaunch", "-optimization", "adam", "-testerlast", "-epoch", "1", "-embeddingDim",
                    "8", "-wordHiddenDim", "10", "-numFilters", "16", "-maxnegsample", "5", "-synthTrain", "24",
                    "-synthAnswers", "20", "-synthVocab", "100", "-maxEvals", "2", "-validSleepTime", "0.2",
                    "-outputprefix", prefix], module=True, timeout=300)
    assert rc == 0
    saved = [f for f in os.listdir(tmp_path) if f.endswith("_model")]
    assert saved, "tester wrote no checkpoint"
    from mpit_b200 import t7
    p = t7.load(str(tmp_path / saved[0]))
    assert p.dim() == 1 and p.numel() == 100 * 8 + 10 * 8 + 10 + 16 * 20 + 16
