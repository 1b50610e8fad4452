"""BiCNN answer-selection sentence CNN (reference: /root/reference/BiCNN/bicnn.lua:30-121).

Four weight-tied towers (Q, Q2, A+, A-) collapse into ONE encoder applied to a padded batch of
sentences::

    LookupTable(V,100) -> Linear(100->200) -> Tanh -> TemporalConvolution(200->3000, kw=2)
        -> Max over time -> ReLU -> Normalize(L2)

followed by the GESD similarity ``1/(1+||q-a||) * 1/(1+exp(-(q.a+1)))`` (bicnn.lua:98-105, restated at
:442) and ``MarginRankingCriterion(margin)`` (:121).  The reference runs one unpadded sentence at a time
through nngraph with weight sharing by ``:set`` on weight *and* gradWeight (:34-74); batching all
sentences of a step into one GEMM with a length mask gives the same math and makes the shared-gradient
accumulation a plain sum.

Flat parameter order = ``model:getParameters()`` of the mmode-1 graph: lookup.weight [V,100],
linear.weight [200,100], linear.bias [200], conv.weight [3000, 2*200] (frame-major: kw*200+in),
conv.bias [3000].
"""
from __future__ import annotations

import math
from typing import List, Optional, Tuple

import torch


class NormalizeL2(torch.autograd.Function):
    """nn.Normalize(2) of BiCNN/Normalize.lua:11-38 with its backward in closed form.

    The reference builds the full d x d Jacobian per row (Normalize.lua:53-69, O(d^2) with d=3000);
    algebraically ``J g = g/norm - x (x.g)/(normp*norm)``."""

    @staticmethod
    def forward(ctx, x, eps=1e-10):
        normp = (x * x).sum(dim=1, keepdim=True) + eps
        norm = normp.sqrt()
        ctx.save_for_backward(x, normp, norm)
        return x / norm

    @staticmethod
    def backward(ctx, g):
        x, normp, norm = ctx.saved_tensors
        xg = (x * g).sum(dim=1, keepdim=True)
        return g / norm - x * (xg / (normp * norm)), None


def gesd(q: torch.Tensor, a: torch.Tensor) -> torch.Tensor:
    """GESD similarity of row-paired embeddings (bicnn.lua:87-105): DotProduct, PairwiseDistance(2),
    ``1/(1+l2) * 1/(1+exp(-(dot+1)))``.  PairwiseDistance adds 1e-6 to the difference? (Torch7's does
    not for p=2 in the version the reference targets), so neither do we."""
    dot = (q * a).sum(dim=1)
    l2 = (q - a).norm(dim=1)
    return 1.0 / (1.0 + l2) * (1.0 / (1.0 + torch.exp(-(dot + 1.0))))


def margin_ranking(s_pos: torch.Tensor, s_neg: torch.Tensor, margin: float) -> torch.Tensor:
    """MarginRankingCriterion(margin) with y=1 (bicnn.lua:121,380): max(0, -(s+ - s-) + margin), per sample."""
    return torch.clamp(margin - (s_pos - s_neg), min=0.0)


class BiCNN(torch.nn.Module):
    def __init__(self, vocab: int, embedding_dim: int = 100, word_hidden_dim: int = 200, num_filters: int = 3000,
                 cont_conv_width: int = 2, embeddings: Optional[torch.Tensor] = None):
        super().__init__()
        self.V, self.E, self.H, self.F, self.kw = vocab, embedding_dim, word_hidden_dim, num_filters, cont_conv_width
        self.lookup = torch.nn.Parameter(torch.empty(vocab, embedding_dim))
        self.lin_w = torch.nn.Parameter(torch.empty(word_hidden_dim, embedding_dim))
        self.lin_b = torch.nn.Parameter(torch.empty(word_hidden_dim))
        self.conv_w = torch.nn.Parameter(torch.empty(num_filters, cont_conv_width * word_hidden_dim))
        self.conv_b = torch.nn.Parameter(torch.empty(num_filters))
        with torch.no_grad():
            if embeddings is not None:
                self.lookup.copy_(embeddings)          # pretrained word vectors (bicnn.lua:34)
            else:
                self.lookup.normal_(0, 1)              # nn.LookupTable default
            for w, b, fan in ((self.lin_w, self.lin_b, embedding_dim),
                              (self.conv_w, self.conv_b, cont_conv_width * word_hidden_dim)):
                s = 1.0 / math.sqrt(fan)
                w.uniform_(-s, s)
                b.uniform_(-s, s)

    def param_list(self):
        return [self.lookup, self.lin_w, self.lin_b, self.conv_w, self.conv_b]

    def encode(self, tokens: torch.Tensor, lens: torch.Tensor) -> torch.Tensor:
        """tokens [S, L] (0-based ids, padded), lens [S] -> L2-normalised sentence embeddings [S, F]."""
        S, L = tokens.shape
        emb = self.lookup[tokens]                                        # [S, L, E]
        h = torch.tanh(emb @ self.lin_w.t() + self.lin_b)                # [S, L, H]
        frames = torch.cat([h[:, i:L - self.kw + 1 + i] for i in range(self.kw)], dim=2)   # [S, L-kw+1, kw*H]
        c = frames @ self.conv_w.t() + self.conv_b                       # [S, T, F]
        T = c.shape[1]
        valid = torch.arange(T, device=tokens.device)[None, :] < (lens[:, None] - self.kw + 1)
        c = c.masked_fill(~valid[:, :, None], float("-inf"))
        m = c.max(dim=1).values
        return NormalizeL2.apply(torch.relu(m))

    def forward(self, q_tokens, q_lens, a_tokens, a_lens):
        return gesd(self.encode(q_tokens, q_lens), self.encode(a_tokens, a_lens))


def pad_sentences(sents: List[torch.Tensor], device=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """List of 1-D id tensors -> (tokens [S, Lmax] padded with 0, lens [S])."""
    lens = torch.tensor([int(s.numel()) for s in sents], dtype=torch.long)
    L = int(lens.max())
    out = torch.zeros(len(sents), L, dtype=torch.long)
    for i, s in enumerate(sents):
        out[i, :s.numel()] = s.long()
    if device is not None:
        out, lens = out.to(device), lens.to(device)
    return out, lens
