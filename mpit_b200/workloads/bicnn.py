"""BiCNN training / evaluation workload (reference: /root/reference/BiCNN/bicnn.lua:279-651,
prepareData.lua, plaunch.lua).

``feval`` (bicnn.lua:279-420) does, per training sample: forward Q and A+, draw up to ``maxnegsample``
random wrong answers and stop at the first one that violates the margin, skip the sample if none does,
back-propagate the hinge loss, add L1/L2 regularisation and clamp the gradient to ``+-gradClip``.  The
reference evaluates the candidates one nngraph forward at a time; here all candidates of all samples of
the mini-batch are encoded in ONE batched pass and the first violator is picked on the device, with the
same random draws in the same order, so the selected triples are the ones the reference would select.
(With ``batchSize > 1`` regularisation and clamping are applied once per mini-batch rather than after
every sample; for the reference's default ``batchSize=1`` the two coincide.)

Corpus files (insuranceQA "ibmprep" layout, prepareData.lua:40-250) are parsed when present; otherwise a
synthetic corpus of the same structure is generated (questions share vocabulary with their answers so
that the task is learnable).
"""
from __future__ import annotations

import os
import random
import time
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import torch

from .. import t7
from ..models.bicnn import BiCNN, gesd, margin_ranking, pad_sentences
from ..utils import Timers

SENTBEGIN, SENTEND = "SENTBEGIN", "SENTEND"


@dataclass
class Corpus:
    word2idx: Dict[str, int]
    idx2word: List[str]
    vectors: torch.Tensor                    # [V, E] initial embeddings (pretrained or random for OOV)
    answers: List[torch.Tensor]              # label -> token ids (mapLabel2AnswerIdx)
    train: List[Tuple[List[int], torch.Tensor, torch.Tensor]]      # (labels, question, answer)
    valid: List[Tuple[List[int], torch.Tensor, List[int]]]         # (targets, question, pool)
    test1: List[Tuple[List[int], torch.Tensor, List[int]]] = field(default_factory=list)
    test2: List[Tuple[List[int], torch.Tensor, List[int]]] = field(default_factory=list)


def _wrap(ids: List[int], begin: int, end: int, kw: int) -> torch.Tensor:
    """prepareData.lua:90,102: ``contConvWidth`` SENTBEGIN tokens, the words, ``contConvWidth-1`` SENTEND."""
    return torch.tensor([begin] * kw + ids + [end] * (kw - 1), dtype=torch.long)


def synthetic_corpus(vocab: int = 2000, n_answers: int = 200, n_train: int = 400, n_eval: int = 60, pool: int = 20,
                     emb_dim: int = 100, kw: int = 2, seed: int = 0) -> Corpus:
    rng = random.Random(seed)
    words = [SENTBEGIN, SENTEND] + [f"w{i}" for i in range(vocab - 2)]
    w2i = {w: i for i, w in enumerate(words)}
    g = torch.Generator().manual_seed(seed)
    vectors = torch.rand(vocab, emb_dim, generator=g)     # math.random() OOV embeddings (prepareData.lua:95-97)
    topic_words = [[rng.randrange(2, vocab) for _ in range(12)] for _ in range(n_answers)]
    answers = [_wrap([rng.choice(tw) for _ in range(rng.randint(6, 14))], 0, 1, kw) for tw in topic_words]

    def question(label):
        return _wrap([rng.choice(topic_words[label]) for _ in range(rng.randint(4, 9))], 0, 1, kw)

    train = []
    for _ in range(n_train):
        lab = rng.randrange(n_answers)
        train.append(([lab], question(lab), answers[lab]))

    def evalset(n):
        out = []
        for _ in range(n):
            lab = rng.randrange(n_answers)
            cand = rng.sample(range(n_answers), pool - 1)
            if lab not in cand:
                cand.append(lab)
            out.append(([lab], question(lab), cand))
        return out

    return Corpus(w2i, words, vectors, answers, train, evalset(n_eval), evalset(n_eval), evalset(n_eval))


def parse_corpus(opt: dict) -> Corpus:
    """prepareData.lua restated: embeddings file, label2answers, train/valid/test files."""
    kw, E = opt["contConvWidth"], opt["embeddingDim"]
    w2i: Dict[str, int] = {}
    i2w: List[str] = []
    vecs: List[List[float]] = []

    def add(word, vec=None):
        if word not in w2i:
            w2i[word] = len(i2w)
            i2w.append(word)
            vecs.append(vec if vec is not None else [random.random() for _ in range(E)])
        return w2i[word]

    with open(opt["embeddingFile"]) as f:
        for line in f:
            parts = line.split()
            if len(parts) == E + 1:
                add(parts[0], [float(x) for x in parts[1:]])
    add(SENTBEGIN)
    add(SENTEND)
    ids = lambda text: _wrap([add(w) for w in text.split()], w2i[SENTBEGIN], w2i[SENTEND], kw)
    answers: Dict[int, torch.Tensor] = {}
    with open(opt["label2answFile"]) as f:
        for line in f:
            lab, text = line.rstrip("\n").split("\t", 1)
            answers[int(lab)] = ids(text)
    n_lab = max(answers) if answers else 0
    ans_list = [answers.get(i + 1, torch.zeros(kw * 2 - 1, dtype=torch.long)) for i in range(n_lab)]   # 1-based labels

    train = []
    with open(opt["trainFile"]) as f:
        for line in f:
            p = line.rstrip("\n").split("\t")
            if len(p) >= 4:
                train.append(([int(x) - 1 for x in p[0].split()], ids(p[2]), ids(p[3])))

    def evalfile(path):
        out = []
        if not path or not os.path.exists(path):
            return out
        with open(path) as f:
            for line in f:
                p = line.rstrip("\n").split("\t")
                if len(p) >= 3:
                    out.append(([int(x) - 1 for x in p[0].split()], ids(p[1]), [int(x) - 1 for x in p[2].split()]))
        return out

    return Corpus(w2i, i2w, torch.tensor(vecs, dtype=torch.float32), ans_list, train, evalfile(opt["validFile"]),
                  evalfile(opt.get("testFile1")), evalfile(opt.get("testFile2")))


def save_binary(corpus: Corpus, prefix: str = "binary_") -> None:
    """The ``-preloadBinary`` cache (plaunch.lua:218-229) in Torch7 format (1-based ids, like the reference)."""
    t7.save(prefix + "mapWordStr2WordIdx", {w: i + 1 for w, i in corpus.word2idx.items()})
    t7.save(prefix + "mapWordIdx2WordStr", {i + 1: w for i, w in enumerate(corpus.idx2word)})
    t7.save(prefix + "mapWordIdx2Vector", {i + 1: v.tolist() for i, v in enumerate(corpus.vectors)})
    t7.save(prefix + "mapLabel2AnswerIdx", {i + 1: (a + 1).float() for i, a in enumerate(corpus.answers)})


class BiCNNTrainer:
    """feval / test / driver around a :class:`BiCNN` with flat parameters (bicnn.lua:245-262)."""

    def __init__(self, corpus: Corpus, opt: dict, device="cpu", rank: int = 0):
        self.c, self.opt, self.rank = corpus, opt, rank
        self.dev = torch.device(device)
        self.model = BiCNN(len(corpus.idx2word), opt["embeddingDim"], opt["wordHiddenDim"], opt["numFilters"],
                           opt["contConvWidth"], corpus.vectors).to(self.dev)
        ps = self.model.param_list()
        n = sum(p.numel() for p in ps)
        self.parameters = torch.zeros(n, device=self.dev)
        self.gradParameters = torch.zeros(n, device=self.dev)
        o = 0
        for p in ps:                      # model:getParameters(): one flat storage the tensors alias
            k = p.numel()
            self.parameters[o:o + k].copy_(p.detach().reshape(-1))
            p.data = self.parameters[o:o + k].view_as(p)
            p.grad = self.gradParameters[o:o + k].view_as(p)
            o += k
        if opt.get("loadmodel", "none") != "none":
            self.parameters.copy_(t7.load(opt["loadmodel"]).to(self.dev))     # bicnn.lua:259-261
        self.rng = random.Random(rank)    # math.randomseed(rank) (plaunch.lua:114-115)
        self.tm = Timers()
        self.loss, self.loss_times = 0.0, 0
        self.best = {}
        self.t_start = time.perf_counter()
        self.batch: List[int] = []
        self._ans_tokens, self._ans_lens = pad_sentences(corpus.answers, self.dev)

    # -- feval (bicnn.lua:279-420) ----------------------------------------------------------------
    def _draw_negatives(self, targets: List[int]) -> List[int]:
        """The reference's rejection sampling (bicnn.lua:327-334), all ``maxnegsample`` draws up front."""
        tset, A = set(targets), len(self.c.answers)
        out = []
        for _ in range(self.opt["maxnegsample"]):
            while True:
                j = self.rng.randrange(A)
                if j not in tset:
                    break
            out.append(j)
        return out

    def feval(self, x: torch.Tensor):
        opt, m, dev = self.opt, self.model, self.dev
        if x is not self.parameters:
            self.parameters.copy_(x)
        self.gradParameters.zero_()
        samples = [self.c.train[i] for i in self.batch]
        K = opt["maxnegsample"]
        with self.tm("fprop"), torch.no_grad():
            qt, ql = pad_sentences([s[1] for s in samples], dev)
            at, al = pad_sentences([s[2] for s in samples], dev)
            negs = [self._draw_negatives(s[0]) for s in samples]
            q = m.encode(qt, ql)
            s_pos = gesd(q, m.encode(at, al))
            neg_idx = torch.tensor(negs, device=dev)                           # [B, K]
            cand = m.encode(self._ans_tokens[neg_idx.view(-1)], self._ans_lens[neg_idx.view(-1)])
            s_neg = gesd(q.repeat_interleave(K, 0), cand).view(len(samples), K)
            viol = (s_pos[:, None] - s_neg) < opt["margin"]                    # bicnn.lua:351-358
            first = torch.where(viol.any(1), viol.float().argmax(1), torch.full((len(samples),), -1, device=dev))
        keep = (first >= 0).nonzero().flatten().tolist()
        f = 0.0
        if keep:
            with self.tm("bprop"):
                sel = [negs[i][int(first[i])] for i in keep]
                qt2, ql2 = qt[keep], ql[keep]
                nt, nl = self._ans_tokens[sel], self._ans_lens[sel]
                qv = m.encode(qt2, ql2)
                loss = margin_ranking(gesd(qv, m.encode(at[keep], al[keep])), gesd(qv, m.encode(nt, nl)),
                                      opt["margin"]).sum()
                loss.backward()
                f = float(loss.detach())
            g, p = self.gradParameters, self.parameters
            if opt["L1reg"] != 0:                                              # bicnn.lua:398-402
                f += opt["L1reg"] * float(p.abs().sum())
                g.add_(torch.sign(p) * opt["L1reg"])
            if opt["L2reg"] != 0:                                              # bicnn.lua:403-408
                f += opt["L2reg"] * float((p * p).sum()) / 2
                g.add_(p * opt["L2reg"])
            g.clamp_(-opt["gradClip"], opt["gradClip"])                        # bicnn.lua:409
        self.loss += f
        self.loss_times += 1
        if self.loss_times % 2000 == 0:
            print(f"Client {self.rank}: curr time: {time.perf_counter() - self.t_start + opt['prevtime']:.2f}, "
                  f"training loss avg. : {self.loss / self.loss_times:.5f} ")
            self.loss, self.loss_times = 0.0, 0
        return f, self.gradParameters

    # -- evaluation (bicnn.lua:422-571) --------------------------------------------------------------
    @torch.no_grad()
    def test(self, dataset, name="valid") -> float:
        if not dataset:
            return 0.0
        with self.tm("test"):
            m = self.model
            emb = torch.cat([m.encode(self._ans_tokens[i:i + 256], self._ans_lens[i:i + 256])
                             for i in range(0, len(self.c.answers), 256)])
            qt, ql = pad_sentences([d[1] for d in dataset], self.dev)
            qv = torch.cat([m.encode(qt[i:i + 256], ql[i:i + 256]) for i in range(0, len(dataset), 256)])
            correct = 0
            for i, (targets, _, pool) in enumerate(dataset):
                sims = gesd(qv[i:i + 1].expand(len(pool), -1), emb[pool])
                # ``simi >= most_simi`` keeps the LAST maximum (bicnn.lua:443-446)
                best = len(pool) - 1 - int(torch.flip(sims, [0]).argmax())
                correct += int(pool[best] in targets)
        acc = correct / len(dataset)
        self.best[name] = max(self.best.get(name, 0.0), acc)
        print(f"Client {self.rank}: curr time: {time.perf_counter() - self.t_start + self.opt['prevtime']:.2f}, "
              f"Accuracy: {acc:.4f}, best Accuracy: {self.best[name]:.4f} on {name} ")
        return acc

    def test3(self):
        return [self.test(self.c.valid, "file1"), self.test(self.c.test1, "file2"), self.test(self.c.test2, "file3")]

    def save(self, prefix: str) -> str:
        """bicnn.lua:590-594: the flat parameter tensor as ``<prefix>_<elapsed>_model`` (Torch7)."""
        path = f"{prefix}_{time.perf_counter() - self.t_start + self.opt['prevtime']:010.2f}_model"
        t7.save(path, self.parameters.detach().cpu())
        return path


def optimizer_for(opt: dict, pc):
    """bicnn.lua:126-245: option table -> (optimizer function, config)."""
    from ..optim import reference as R
    name = opt["optimization"]
    if name == "sgd":
        return R.msgdsingle, {"lr": opt["learningRate"], "lrd": opt["weightDecay"], "mom": opt["momentum"], "pclient": pc}
    if name == "downpour":
        return R.downpour, {"lr": opt["learningRate"], "lrd": opt["weightDecay"], "pclient": pc, "su": opt["commperiod"]}
    if name == "eamsgd":
        return R.eamsgd, {"lr": opt["learningRate"], "lrd": opt["weightDecay"], "pclient": pc, "su": opt["commperiod"],
                          "mva": opt["movingrate"], "mom": opt["momentum"]}
    rules = {
        "rmsprop": dict(mode=opt["modeRMSProp"], decay=opt["decayRMSProp"], lr=opt["lrRMSProp"],
                        momentum=opt["momentumRMSProp"], epsilon=opt["epsilonRMSProp"]),
        "adam": dict(mode=opt["modeAdam"], lr=opt["lrAdam"], beta1=opt["beta1Adam"], beta2=opt["beta2Adam"],
                     epsilon=opt["epsilonAdam"]),
        "adagrad": dict(mode=opt["modeAdagrad"], lr=opt["lrAdagrad"], lrd=opt["lrDecayAdagrad"],
                        epsilon=opt["epsilonAdagrad"]),
        "adadelta": dict(mode=opt["modeAdadelta"], rho=opt["rhoAdadelta"], lr=opt["lrAdadelta"],
                         epsilon=opt["epsilonAdadelta"]),
    }
    rules["adamax"] = dict(rules["adam"])
    base = name[:-6] if name.endswith("single") else name
    if base not in rules:
        raise ValueError("unknown optimization method")
    cfg = dict(rules[base])
    cfg["pclient"] = pc
    if not name.endswith("single"):
        cfg["su"] = opt["commperiod"]
    return R.OPTIMIZERS[name], cfg


def run_trainer(tr: BiCNNTrainer, pc, is_tester: bool, also_test: bool) -> dict:
    """bicnn.lua:575-651: the tester role loops pull -> test3 -> save -> sleep (bounded by ``maxEvals``; the
    reference never exits, TODO at :581); trainers run epochs of mini-batches through the optimizer."""
    opt = tr.opt
    if pc is not None:
        pc.start(tr.parameters, tr.gradParameters)
    out = {"acc": [], "loss": []}
    if is_tester:
        for _ in range(opt.get("maxEvals", 3)):
            t0 = time.perf_counter()
            pc.async_recv_param()
            pc.wait()
            print(f"Client {tr.rank}: communication time: {time.perf_counter() - t0:.2f} ")
            out["acc"].append(tr.test3())
            if opt["outputprefix"] != "none":
                tr.save(opt["outputprefix"])
            time.sleep(opt["validSleepTime"])
    else:
        opti, cfg = optimizer_for(opt, pc)
        n = len(tr.c.train)
        pversion = 0
        for epoch in range(1, opt["epoch"] + 1):
            t0 = time.perf_counter()
            shuffle = torch.randperm(n).tolist()
            for t in range(0, n, opt["batchSize"]):
                tr.batch = shuffle[t:t + opt["batchSize"]]
                _, fx = opti(tr.feval, tr.parameters, cfg)
                out["loss"].append(fx[0])
                if also_test and pversion % opt["commperiod"] == 0 and pversion % opt.get("testEvery", 50) == 0:
                    out["acc"].append(tr.test3())
                    if opt["outputprefix"] != "none":
                        tr.save(opt["outputprefix"])
                pversion += 1
            print(f"client {tr.rank}: epoch {epoch} done,  for {time.perf_counter() - t0} seconds ")
        tr.tm["sync"] = cfg.get("dusync", 0.0)
    print("Before Stop !!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!")
    if pc is not None:
        pc.stop()
    print(tr.rank, "total training time", time.perf_counter() - tr.t_start)
    print(tr.rank, "factor time", dict(tr.tm))
    return out


class BiCNNKernelTrainer(BiCNNTrainer):
    """``-type cuda``: the same feval / test semantics on one B200 through the kernel engine
    (models/bicnn_engine.py).  One encoder pass covers the question, the positive answer and all
    ``maxnegsample`` candidate negatives of every sample of the mini-batch; the first margin violator is
    picked on the device from that pass and the SAME pass is back-propagated for the selected triples
    (unselected sentences simply receive a zero gradient), so no second forward is needed."""

    def __init__(self, corpus: Corpus, opt: dict, device="cuda", rank: int = 0):
        super().__init__(corpus, opt, device, rank)
        from ..models.bicnn_engine import BiCnnEngine
        from ..ops import grad_reg_clip
        self._reg_clip = grad_reg_clip
        K, B = opt["maxnegsample"], opt["batchSize"]
        lmax = max(int(a.numel()) for a in corpus.answers)
        lmax = max([lmax] + [int(s[1].numel()) for s in corpus.train] + [int(s[1].numel()) for s in corpus.valid])
        self.Lk = (lmax + 31) // 32 * 32
        self.S_train = (B * (2 + K) + 3) // 4 * 4
        mk = lambda S: BiCnnEngine(len(corpus.idx2word), S, self.Lk, opt["embeddingDim"], opt["wordHiddenDim"],
                                   opt["numFilters"], opt["contConvWidth"], self.dev, opt["margin"])
        self.eng = mk(self.S_train)
        self.eng_eval = mk(256)
        self._ans_tok32 = self._fit(self._ans_tokens)

    def _fit(self, tok: torch.Tensor) -> torch.Tensor:
        out = torch.zeros(tok.shape[0], self.Lk, dtype=torch.int32, device=self.dev)
        out[:, :tok.shape[1]] = tok.int()
        return out

    def feval(self, x: torch.Tensor):
        opt, eng, dev = self.opt, self.eng, self.dev
        if x is not self.parameters:
            self.parameters.copy_(x)
        self.gradParameters.zero_()
        samples = [self.c.train[i] for i in self.batch]
        B, K = len(samples), opt["maxnegsample"]
        with self.tm("fprop"):
            qt, ql = pad_sentences([s[1] for s in samples], dev)
            at, al = pad_sentences([s[2] for s in samples], dev)
            negs = torch.tensor([self._draw_negatives(s[0]) for s in samples], device=dev)       # [B, K]
            tok = torch.cat([self._fit(qt), self._fit(at), self._ans_tok32[negs.view(-1)]])
            lens = torch.cat([ql, al, self._ans_lens[negs.view(-1)]]).int()
            eng.load_flat(self.parameters)
            eng.set_sentences(tok, lens)
            eng.encode()
            ar = torch.arange(B, device=dev, dtype=torch.int32)
            s_pos = eng.gesd(ar, ar + B)
            qi_all = ar.repeat_interleave(K)
            ai_all = (2 * B + torch.arange(B * K, device=dev)).int()
            s_neg = eng.gesd(qi_all, ai_all).view(B, K)
            viol = (s_pos[:, None] - s_neg) < opt["margin"]
            has = viol.any(1)
            first = viol.float().argmax(1)
        keep = has.nonzero().flatten()
        f = 0.0
        if keep.numel():
            with self.tm("bprop"):
                k32 = keep.int()
                neg_sent = (2 * B + keep * K + first[keep]).int()
                qi = torch.stack([k32, k32], 1).view(-1).contiguous()
                ai = torch.stack([k32 + B, neg_sent], 1).view(-1).contiguous()
                loss = eng.loss_backward(qi, ai, self.gradParameters)
                eng.flat_grad(self.gradParameters)
                f = float(loss.item())
            p = self.parameters
            if opt["L1reg"] != 0:
                f += opt["L1reg"] * float(p.abs().sum())
            if opt["L2reg"] != 0:
                f += opt["L2reg"] * float((p * p).sum()) / 2
            # L1 / L2 / clamp in one fused pass over the flat gradient (bicnn.lua:398-409)
            self._reg_clip(self.gradParameters, p, l1=opt["L1reg"], l2=opt["L2reg"], clip=opt["gradClip"])
        self.loss += f
        self.loss_times += 1
        return f, self.gradParameters

    @torch.no_grad()
    def _encode_all(self, tok: torch.Tensor, lens: torch.Tensor) -> torch.Tensor:
        eng = self.eng_eval
        eng.load_flat(self.parameters)
        out = []
        for i in range(0, tok.shape[0], eng.S):
            eng.set_sentences(tok[i:i + eng.S], lens[i:i + eng.S].int())
            out.append(eng.encode()[:min(eng.S, tok.shape[0] - i)].clone())
        return torch.cat(out)

    @torch.no_grad()
    def test(self, dataset, name="valid") -> float:
        if not dataset:
            return 0.0
        with self.tm("test"):
            emb = self._encode_all(self._ans_tok32, self._ans_lens)
            qt, ql = pad_sentences([d[1] for d in dataset], self.dev)
            qv = self._encode_all(self._fit(qt), ql)
            correct = 0
            for i, (targets, _, pool) in enumerate(dataset):
                sims = gesd(qv[i:i + 1].expand(len(pool), -1), emb[pool])
                best = len(pool) - 1 - int(torch.flip(sims, [0]).argmax())
                correct += int(pool[best] in targets)
        acc = correct / len(dataset)
        self.best[name] = max(self.best.get(name, 0.0), acc)
        print(f"Client {self.rank}: curr time: {time.perf_counter() - self.t_start + self.opt['prevtime']:.2f}, "
              f"Accuracy: {acc:.4f}, best Accuracy: {self.best[name]:.4f} on {name} ")
        return acc


def device_rule_for(opt: dict):
    """Option table -> (rule, hyper, gscale) of the asynchronous device parameter server: the server-side rules of
    BiCNN/pserver.lua:123-198 ('global' modes) and Downpour's plain add of ``-lr * g`` (optim-downpour.lua:24-54)."""
    name = opt["optimization"]
    if name == "downpour":
        return "add", {}, -float(opt["learningRate"])
    if name == "rmsprop":
        if opt["modeRMSProp"] != "global":
            raise ValueError("device path: rmsprop runs in 'global' mode (the rule lives with the shard)")
        return "rmsprop", dict(decay=opt["decayRMSProp"], lr=opt["lrRMSProp"], momentum=opt["momentumRMSProp"],
                               epsilon=opt["epsilonRMSProp"]), 1.0
    if name in ("adam", "adamax"):
        h = dict(lr=opt["lrAdam"], beta1=opt["beta1Adam"], beta2=opt["beta2Adam"], epsilon=opt["epsilonAdam"])
        if name == "adam":
            h["step_div"] = opt["stepDivAdam"]            # pserver.lua:151-152
        return name, h, 1.0
    if name == "adagrad":
        return "adagrad", dict(lr=opt["lrAdagrad"], lrd=opt["lrDecayAdagrad"], epsilon=opt["epsilonAdagrad"]), 1.0
    if name == "adadelta":
        return "adadelta", dict(rho=opt["rhoAdadelta"], lr=opt["lrAdadelta"], epsilon=opt["epsilonAdadelta"]), 1.0
    raise ValueError(f"device path: optimization {name!r} is not a server-side rule (use -type float)")


class BiCNNDeviceTrainer(BiCNNKernelTrainer):
    """Distributed BiCNN training on the NVLink device path (BASELINE.json config 4): one process per GPU, every GPU
    a trainer and the owner of one shard of the parameters with the server-side rule state
    (``optim.DeviceRuleParallel`` over ``parallel.AsyncDevicePS``).

    A training step never returns to the host: the step's sample indices are copied in, and one CUDA graph runs the
    device-side batch assembly with negative sampling (bicnn.lua:321-334), the encoder over question / positive /
    candidates, the first-violator selection (:351-358), the backward pass, the client's push (L1 / L2 / clamp of
    :398-409, ``commperiod`` accumulation), the owner's drain and the parameter pull.  Losses accumulate in device
    memory and are read when somebody asks (:412-418 prints an average every 2000 evaluations)."""

    def __init__(self, corpus: Corpus, opt: dict, device, rank: int, heap, server: str = "thread",
                 use_graphs: bool = True):
        super().__init__(corpus, opt, device, rank)
        from ..optim.device import DeviceRuleParallel
        from ..parallel.center import ShardedCenter
        dev, L = self.dev, self.Lk
        B, K = opt["batchSize"], opt["maxnegsample"]
        fit = lambda sents: self._fit(pad_sentences(sents, dev)[0])
        lens = lambda sents: torch.tensor([int(s.numel()) for s in sents], dtype=torch.int32, device=dev)
        qs, as_ = [s[1] for s in corpus.train], [s[2] for s in corpus.train]
        tmax = max(1, max(len(s[0]) for s in corpus.train))
        targets = torch.full((len(corpus.train), tmax), -1, dtype=torch.int32)
        for i, s in enumerate(corpus.train):
            targets[i, :len(s[0])] = torch.tensor(s[0], dtype=torch.int32)
        self.eng.attach_dataset(fit(qs), lens(qs), fit(as_), lens(as_), targets.to(dev), self._ans_tok32,
                                self._ans_lens.int().contiguous(), B, K, seed=1000003 * (rank + 1))
        rule, hyper, gscale = device_rule_for(opt)
        self.center = ShardedCenter.create(heap, self.parameters.numel())
        self.optim = DeviceRuleParallel(self.eng, self.parameters, rule, hyper, self.center, heap, su=opt["commperiod"],
                                        l1=opt["L1reg"], l2=opt["L2reg"], clip=opt["gradClip"], gscale=gscale,
                                        server=server, use_graphs=use_graphs, seed_center=(rank == 0))
        self.parameters = self.optim.x                      # the flat vector the step works on
        self.reg_loss = torch.zeros(2, device=dev)
        # The step's sample indices reach the device by an asynchronous copy from pinned memory, and with the step in a
        # CUDA graph the host runs many steps ahead of the GPU: every step in flight needs its own pinned slot (one
        # buffer refilled per step would be read by the copy engine AFTER later steps had overwritten it).
        self._batch_ring = [torch.full((B,), -1, dtype=torch.int32).pin_memory() for _ in range(32)]
        self._batch_copied = [torch.cuda.Event() for _ in self._batch_ring]
        self._staged = 0
        self.steps = 0

    def _stage_batch(self, batch: List[int]) -> None:
        """Sample indices of the next step -> ``eng.batch`` (stream-ordered copy from this step's pinned slot)."""
        slot = self._staged % len(self._batch_ring)
        if self._staged >= len(self._batch_ring):
            self._batch_copied[slot].synchronize()          # only blocks when the GPU is a whole ring behind
        h = self._batch_ring[slot]
        h.fill_(-1)
        h[:len(batch)] = torch.tensor(batch, dtype=torch.int32)
        self.eng.batch.copy_(h, non_blocking=True)
        self._batch_copied[slot].record()
        self._staged += 1

    def train_step(self, batch: List[int]) -> None:
        """One mini-batch: ``batch`` lists sample indices (at most ``batchSize``)."""
        o = self.optim
        self._stage_batch(batch)
        do_sync = o.t % o.su == 0

        def body():
            self.eng.train_fwd_bwd()
            o.exchange(do_sync, active=self.eng.active, reg_loss=self.reg_loss)

        o.graphs.run(("bicnn", do_sync), body)
        o.t += 1
        self.steps += 1

    def loss_stats(self) -> dict:
        """Sums since the start (one device->host read): hinge loss, regularisation terms, triples, steps."""
        c = self.eng.counters.tolist()
        return {"hinge": float(self.eng.loss_acc.item()), "l1": float(self.reg_loss[0]), "l2": float(self.reg_loss[1]),
                "triples": int(c[0]), "steps": int(c[1])}

    def pull_params(self) -> torch.Tensor:
        """Tester role (bicnn.lua:580-596): fetch the current parameters from the shard owners."""
        self.optim.ps.pull(self.optim.x, self.eng.wpk, self.eng.wmap, self.eng.map_base)
        torch.cuda.synchronize(self.dev)
        return self.optim.x

    def close(self) -> int:
        return self.optim.close()


def run_device_trainer(tr: BiCNNDeviceTrainer, is_tester: bool, also_test: bool, barrier=None) -> dict:
    """bicnn.lua:575-651 on the device path: the tester pulls / evaluates / saves (bounded by ``maxEvals``), trainers
    run epochs of mini-batches; nobody stops serving its shard before everybody is done."""
    opt = tr.opt
    out = {"acc": [], "loss": [], "steps": 0, "seconds": 0.0}
    t_run = time.perf_counter()
    if is_tester:
        for _ in range(opt.get("maxEvals", 3)):
            t0 = time.perf_counter()
            tr.pull_params()
            print(f"Client {tr.rank}: communication time: {time.perf_counter() - t0:.4f} ")
            out["acc"].append(tr.test3())
            if opt["outputprefix"] != "none":
                tr.save(opt["outputprefix"])
            time.sleep(opt["validSleepTime"])
    else:
        n = len(tr.c.train)
        B = opt["batchSize"]
        prev = tr.loss_stats()
        for epoch in range(1, opt["epoch"] + 1):
            t0 = time.perf_counter()
            shuffle = torch.randperm(n).tolist()
            for t in range(0, n, B):
                tr.train_step(shuffle[t:t + B])
                if also_test and tr.steps % opt.get("testEvery", 50) == 0:
                    tr.pull_params()
                    out["acc"].append(tr.test3())
            torch.cuda.synchronize(tr.dev)
            cur = tr.loss_stats()
            steps = max(1, cur["steps"] - prev["steps"])
            f = (cur["hinge"] - prev["hinge"] + cur["l1"] - prev["l1"] + cur["l2"] - prev["l2"]) / steps
            out["loss"].append(f)
            prev = cur
            print(f"client {tr.rank}: epoch {epoch} done,  for {time.perf_counter() - t0} seconds , "
                  f"training loss avg. : {f:.5f} ")
        out["steps"] = tr.steps
    out["seconds"] = time.perf_counter() - t_run
    print("Before Stop !!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!")
    torch.cuda.synchronize(tr.dev)
    if barrier is not None:
        barrier()                     # every client's last push is out before any owner stops serving
    out["drains"] = tr.close()
    tr.optim.ps.check_timeout()
    print(tr.rank, "total training time", time.perf_counter() - tr.t_start)
    return out
