"""BiCNN encoder + GESD/margin loss on one B200 through mpit_b200 kernels (reference model:
/root/reference/BiCNN/bicnn.lua:30-121).

Every sentence of a step (questions, positive answers, sampled negatives) is one row block of a
padded ``[S, L]`` token matrix; the tower runs once over all of them:

    embedding gather (bicnn.cu) -> Linear+Tanh (tcgen05 igemm, fused bias/tanh epilogue)
      -> TemporalConvolution(kw=2) as an implicit GEMM whose epilogue does bias + frame masking +
         Max-over-time + ReLU with the arg-max (umma_igemm.cu mode MAXTIME)
      -> Normalize(2) (bicnn.cu) -> GESD pairs -> margin ranking

and backward: fused GESD/margin gradient -> closed-form Normalize/ReLU/Max backward into a dense frame
gradient -> tcgen05 weight- and data-gradient kernels -> tied embedding scatter-add straight into the
flat gradient.  Flat parameter order is the reference's (lookup, linear.w, linear.b, conv.w, conv.b).
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from .._native import kernels, check, ptr, cur_stream
from ..ops import local_step, map_copy, colsum
from ..ops.umma_ops import make_tmap, igemm_geom, IgemmLaunch, weight_tmap, EPI_STORE, EPI_MASK

EPI_MAXTIME = 4
MN_SWIZZLE = 2


def _pad(n: int, a: int) -> int:
    return (n + a - 1) // a * a


class BiCnnEngine:
    def __init__(self, vocab: int, S: int, L: int = 32, emb: int = 100, hidden: int = 200, filters: int = 3000,
                 kw: int = 2, device="cuda", margin: float = 0.02):
        assert kw == 2 and L % 32 == 0 and emb % 4 == 0 and S % 4 == 0
        self.V, self.S, self.L, self.E, self.H, self.F, self.kw, self.margin = vocab, S, L, emb, hidden, filters, kw, margin
        self.dev = torch.device(device)
        dev = self.dev
        # Operand paddings.  Wide layers are padded to 128 so that the big GEMMs run on 128-column tiles: with
        # 32/64-column tiles every n-tile re-fetches the whole A operand (the temporal convolution's data gradient
        # re-read 8192 x 6016 floats seven times: 187 us; forward 47 times at 64 columns) -- the 14 % extra K / N of
        # the padding (200 -> 256, 3000 -> 3072) costs far less than that.
        wide = lambda n: _pad(n, 128) if n >= 128 else _pad(n, 32)
        self.Ep, self.Hp, self.Fp = _pad(emb, 32), wide(hidden), (_pad(filters, 128) if filters >= 128 else _pad(filters, 64))
        Ep, Hp, Fp = self.Ep, self.Hp, self.Fp
        self.k = kernels()
        # ---- flat <-> packed layout -------------------------------------------------------------
        self.n_lookup = vocab * emb
        self.n_flat = self.n_lookup + hidden * emb + hidden + filters * kw * hidden + filters
        o = 0
        self.off = {}
        for name, size in (("W1", Hp * Ep), ("b1", Hp), ("W2", Fp * kw * Hp), ("b2", Fp), ("Wd2", Hp * kw * Fp), ("Wd1", Ep * Hp)):
            self.off[name] = o
            o += _pad(size, 32)
        self.n_packed = o
        self.wpk = torch.zeros(o, device=dev)
        self.gpk = torch.zeros(o, device=dev)
        n_dense = self.n_flat - self.n_lookup
        m1 = torch.full((n_dense,), -1, dtype=torch.int64)
        m2 = torch.full((n_dense,), -1, dtype=torch.int64)
        p = 0
        h_i, e_i = torch.meshgrid(torch.arange(hidden), torch.arange(emb), indexing="ij")
        m1[p:p + hidden * emb] = (self.off["W1"] + h_i * Ep + e_i).reshape(-1)
        m2[p:p + hidden * emb] = (self.off["Wd1"] + e_i * Hp + h_i).reshape(-1)
        p += hidden * emb
        m1[p:p + hidden] = self.off["b1"] + torch.arange(hidden)
        p += hidden
        f_i, k_i, hh = torch.meshgrid(torch.arange(filters), torch.arange(kw), torch.arange(hidden), indexing="ij")
        m1[p:p + filters * kw * hidden] = (self.off["W2"] + f_i * (kw * Hp) + k_i * Hp + hh).reshape(-1)
        m2[p:p + filters * kw * hidden] = (self.off["Wd2"] + hh * (kw * Fp) + k_i * Fp + f_i).reshape(-1)
        p += filters * kw * hidden
        m1[p:p + filters] = self.off["b2"] + torch.arange(filters)
        self.wmap = torch.cat([m1, m2]).to(torch.int32).to(dev)     # dense (non-embedding) parameters only
        self.gmap = self.wmap[:n_dense]
        # ---- activations ------------------------------------------------------------------------
        R = S * L
        f32 = dict(dtype=torch.float32, device=dev)
        self.tokens = torch.zeros(S, L, dtype=torch.int32, device=dev)
        self.lens = torch.ones(S, dtype=torch.int32, device=dev)
        self.frames = torch.zeros(S, dtype=torch.int32, device=dev)
        self.row_valid = torch.zeros(R, dtype=torch.int32, device=dev)
        self.emb = torch.zeros(R * Ep, **f32)
        self.h = torch.zeros((R + 32) * Hp, **f32)
        self.mx = torch.zeros(S * Fp, dtype=torch.int64, device=dev)
        self.r = torch.zeros(S * Fp, **f32)
        self.y = torch.zeros(S * Fp, **f32)
        self.arg = torch.zeros(S * Fp, dtype=torch.uint8, device=dev)
        self.normp = torch.zeros(S, **f32)
        self.dy = torch.zeros(S * Fp, **f32)
        self.gmax = torch.zeros(S * Fp, **f32)     # per-sentence gradient w.r.t. the max-over-time output
        self.dC = torch.zeros((R + 32) * Fp, **f32)
        self.dZ1 = torch.zeros(R * Hp, **f32)
        self.dEmb = torch.zeros(R * Ep, **f32)
        self.loss_sum = torch.zeros(1, **f32)
        self.err = torch.zeros(1, dtype=torch.int32, device=dev)
        # flat-order gradient buffer (its embedding part is what the push of a device optimizer reads; the dense
        # part arrives packed in gpk and is gathered through gmap from map_base on)
        self.gflat = torch.zeros(self.n_flat, **f32)
        self.map_base = self.n_lookup
        self._build_plans()

    def _w(self, name, size):
        return self.wpk[self.off[name]:self.off[name] + size]

    def _g(self, name, size):
        return self.gpk[self.off[name]:self.off[name] + size]

    def _build_plans(self):
        S, L, Ep, Hp, Fp, kw = self.S, self.L, self.Ep, self.Hp, self.Fp, self.kw
        R = S * L
        tw_tiles = L // 32
        rows_map = lambda t, C, box_n, sw=1: make_tmap(t.data_ptr(), [C, 1, 1, R], [C * 4, C * 4, C * 4], [32, 1, 1, box_n], sw)
        seq_map = lambda t, C, box_n, sw=1: make_tmap(t.data_ptr(), [C, L, 1, S], [C * 4, L * C * 4, L * C * 4],
                                                      [32, 32, 1, box_n], sw)
        from ..ops.umma_ops import _pick_block
        bn_f = _pick_block(Fp, (128, 64))          # temporal convolution forward / weight gradient tiles
        bn_h = _pick_block(Hp, (128, 64, 32))      # its data gradient (N = hidden)
        # Linear + Tanh
        self.p_lin = IgemmLaunch(
            rows_map(self.emb, Ep, 128), weight_tmap(self._w("W1", Hp * Ep), Hp, Ep, 32),
            igemm_geom(nsub=1, tw=1, th=1, tn=128, tiles_w=1, tiles_h=1, kw_it=1, cb=Ep // 32, num_kb=Ep // 32, sgn=1,
                       B=R, OH=1, OW=1, N=Hp, mode=EPI_STORE, relu=2),
            1.0, self._w("b1", Hp), self.h, None, None, None, None, None, 32)
        # TemporalConvolution + Max over time + ReLU
        self.p_conv = IgemmLaunch(
            seq_map(self.h, Hp, 4), weight_tmap(self._w("W2", Fp * kw * Hp), Fp, kw * Hp, bn_f),
            igemm_geom(nsub=1, tw=32, th=1, tn=4, tiles_w=tw_tiles, tiles_h=1, kw_it=kw, cb=Hp // 32,
                       num_kb=kw * Hp // 32, sgn=1, B=S, OH=1, OW=L, N=Fp, mode=EPI_MAXTIME),
            1.0, self._w("b2", Fp), self.mx, None, None, None, self.frames, None, bn_f)
        # data gradient of the temporal convolution, tanh' fused: dZ1 = (dC * Wd2) .* (1 - h^2)
        self.p_conv_dgrad = IgemmLaunch(
            seq_map(self.dC, Fp, 4), weight_tmap(self._w("Wd2", Hp * kw * Fp), Hp, kw * Fp, bn_h),
            igemm_geom(nsub=1, tw=32, th=1, tn=4, tiles_w=tw_tiles, tiles_h=1, kw_it=kw, cb=Fp // 32,
                       num_kb=kw * Fp // 32, sgn=-1, B=S, OH=1, OW=L, N=Hp, mode=EPI_MASK, relu=2),
            1.0, None, self.dZ1, None, self.h, None, None, None, bn_h)
        # data gradient of the Linear layer: dEmb = dZ1 * Wd1
        self.p_lin_dgrad = IgemmLaunch(
            rows_map(self.dZ1, Hp, 128), weight_tmap(self._w("Wd1", Ep * Hp), Ep, Hp, 32),
            igemm_geom(nsub=1, tw=1, th=1, tn=128, tiles_w=1, tiles_h=1, kw_it=1, cb=Hp // 32, num_kb=Hp // 32, sgn=-1,
                       B=R, OH=1, OW=1, N=Ep, mode=EPI_STORE),
            1.0, None, self.dEmb, None, None, None, None, None, 32)
        # weight gradients (MN-major operands: 32-byte-atom swizzle)
        G = lambda *a: (ctypes.c_int * 11)(*a)
        ceil = lambda a, b: (a + b - 1) // b
        nkb_c = ceil(S, 2) * tw_tiles
        self.w_conv = dict(tx=seq_map(self.h, Hp, 2, MN_SWIZZLE), tz=seq_map(self.dC, Fp, 2, MN_SWIZZLE),
                           geom=G(32, 1, 2, tw_tiles, 1, kw, Hp // 32, kw * Hp // 32, nkb_c, S, Fp), block_n=bn_f,
                           nsplit=min(4, nkb_c), chunks=kw * Hp // 32, N=Fp, gw=self._g("W2", Fp * kw * Hp))
        nkb_l = ceil(R, 64)
        self.w_lin = dict(tx=rows_map(self.emb, Ep, 64, MN_SWIZZLE), tz=rows_map(self.dZ1, Hp, 64, MN_SWIZZLE),
                          geom=G(1, 1, 64, 1, 1, 1, Ep // 32, Ep // 32, nkb_l, R, Hp), block_n=32,
                          nsplit=min(8, nkb_l), chunks=Ep // 32, N=Hp, gw=self._g("W1", Hp * Ep))
        need = 0
        for w in (self.w_conv, self.w_lin):
            w["m_pad"] = ceil(w["chunks"], 4) * 128
            need = max(need, w["nsplit"] * w["m_pad"] * w["N"])
        self.part = torch.empty(need, device=self.dev)

    # -- parameters -----------------------------------------------------------------------------------
    def load_flat(self, theta: torch.Tensor) -> None:
        self.theta = theta
        dense = theta[self.n_lookup:]
        local_step(dense, None, None, None, None, self.wpk, self.wmap, do_apply=False)

    def flat_grad(self, out: torch.Tensor) -> torch.Tensor:
        """Dense-parameter gradients gathered from the packed layout; the embedding gradient was already
        scatter-added straight into ``out[:n_lookup]`` by ``backward``."""
        map_copy(out[self.n_lookup:], self.gpk, self.gmap, scatter=False)
        return out

    def set_sentences(self, tokens: torch.Tensor, lens: torch.Tensor) -> None:
        S, L = self.S, self.L
        n, l = tokens.shape
        assert n <= S and l <= L
        self.tokens.zero_()
        self.tokens[:n, :l].copy_(tokens)
        self.lens.fill_(self.kw)
        self.lens[:n].copy_(lens)
        self.frames.copy_((self.lens - self.kw + 1).clamp(min=0))
        if n < S:
            self.frames[n:] = 0
        self.row_valid.copy_((torch.arange(L, device=self.dev)[None, :] < self.lens[:, None]).view(-1).int())

    # -- forward ------------------------------------------------------------------------------------------
    def encode(self) -> torch.Tensor:
        k, st = self.k, cur_stream
        R = self.S * self.L
        E = self.theta[:self.n_lookup]
        check(k.mpit_embedding_gather(ptr(E), ptr(self.tokens), ptr(self.emb), R, self.E, self.E, self.Ep, st()),
              "mpit_embedding_gather")
        self.p_lin.run(self.err)
        self.mx.zero_()
        self.p_conv.run(self.err)
        check(k.mpit_encoder_finish(ptr(self.mx), ptr(self.r), ptr(self.arg), ptr(self.y), ptr(self.normp), self.S,
                                    self.F, self.Fp, 1e-10, st()), "mpit_encoder_finish")
        return self.y.view(self.S, self.Fp)[:, :self.F]

    def gesd(self, qi: torch.Tensor, ai: torch.Tensor, stats: Optional[torch.Tensor] = None) -> torch.Tensor:
        P = qi.numel()
        sim = torch.empty(P, device=self.dev)
        check(self.k.mpit_gesd_pairs(ptr(self.y), ptr(qi), ptr(ai), ptr(sim), ptr(stats), P, self.F, self.Fp,
                                     cur_stream()), "mpit_gesd_pairs")
        return sim

    # -- loss + backward --------------------------------------------------------------------------------------
    def loss_backward(self, qi: torch.Tensor, ai: torch.Tensor, grad_flat: torch.Tensor) -> torch.Tensor:
        """``qi/ai`` list pairs as (q,a+),(q,a-) per triple.  Fills the packed dense gradients and adds the
        embedding gradient into ``grad_flat[:n_lookup]`` (which the caller has zeroed); returns loss_sum."""
        P = qi.numel()
        stats = torch.empty(2 * P, device=self.dev)
        sim = self.gesd(qi, ai, stats)
        self.loss_sum.zero_()
        return self.backward_from_pairs(qi, ai, sim, stats, grad_flat)

    def backward_from_pairs(self, qi, ai, sim, stats, grad_flat: torch.Tensor, loss_acc: Optional[torch.Tensor] = None):
        """Backward of the margin loss for pair lists whose similarities are already known (``sim``, ``stats`` from
        :meth:`gesd` or the device-side violator selection).  The hinge loss is ADDED to ``loss_acc`` (default: the
        engine's ``loss_sum``)."""
        k, st = self.k, cur_stream
        P = qi.numel()
        if loss_acc is None:
            loss_acc = self.loss_sum
        self.dy.zero_()
        check(k.mpit_gesd_margin_bwd(ptr(self.y), ptr(qi), ptr(ai), ptr(sim), ptr(stats), ptr(self.dy), ptr(loss_acc),
                                     P // 2, self.F, self.Fp, self.margin, st()), "mpit_gesd_margin_bwd")
        # writes every frame of dC (no clearing of the 100 MB tensor) and the per-sentence gradient gmax
        check(k.mpit_encoder_bwd(ptr(self.dy), ptr(self.r), ptr(self.arg), ptr(self.normp), ptr(self.dC), ptr(self.gmax),
                                 self.S, self.F, self.Fp, self.L, st()), "mpit_encoder_bwd")
        R = self.S * self.L
        # conv bias gradient = column sums of the per-sentence gradient (each (sentence, filter) feeds one frame);
        # conv weight gradient on tensor cores
        colsum(self.gmax, self._g("b2", self.Fp), self.S, self.Fp)
        self._wgrad(self.w_conv)
        self.p_conv_dgrad.run(self.err)
        colsum(self.dZ1, self._g("b1", self.Hp), R, self.Hp)
        self._wgrad(self.w_lin)
        self.p_lin_dgrad.run(self.err)
        check(k.mpit_embedding_scatter_add(ptr(grad_flat), ptr(self.tokens), ptr(self.dEmb), ptr(self.row_valid), R,
                                           self.E, self.E, self.Ep, st()), "mpit_embedding_scatter_add")
        return loss_acc

    # -- device-side training step (no host round trip) -----------------------------------------------------------
    def attach_dataset(self, q_tok, q_len, a_tok, a_len, targets, ans_tok, ans_len, B: int, K: int, seed: int) -> None:
        """Device-resident training set for :meth:`train_fwd_bwd`: int32 token matrices padded to ``L`` columns,
        lengths, per-sample target lists (-1 padded) and the answer table.  ``S`` must be >= B*(2+K)."""
        assert self.S >= B * (2 + K)
        i32 = dict(dtype=torch.int32, device=self.dev)
        self.ds = dict(q_tok=q_tok, q_len=q_len, a_tok=a_tok, a_len=a_len, targets=targets, ans_tok=ans_tok,
                       ans_len=ans_len)
        self.B, self.K, self.seed = B, K, seed
        self.batch = torch.full((B,), -1, **i32)
        self.step_ctr = torch.zeros(1, dtype=torch.int64, device=self.dev)
        self.neg_idx = torch.zeros(max(B * K, 1), **i32)
        P = B * (1 + K)
        b = torch.arange(B, **i32)
        self.qi_all = torch.cat([b, b.repeat_interleave(K)]).contiguous()
        self.ai_all = torch.cat([b + B, 2 * B + torch.arange(B * K, **i32)]).contiguous()
        self.sim_all = torch.zeros(P, device=self.dev)
        self.stats_all = torch.zeros(2 * P, device=self.dev)
        self.qi2 = b.repeat_interleave(2).contiguous()
        self.ai2 = torch.zeros(2 * B, **i32)
        self.sim2 = torch.zeros(2 * B, device=self.dev)
        self.stats2 = torch.zeros(4 * B, device=self.dev)
        self.active = torch.zeros(1, **i32)
        self.counters = torch.zeros(4, device=self.dev)      # {triples trained on, steps, -, -}
        self.loss_acc = torch.zeros(1, device=self.dev)      # running sum of the hinge losses
        ptrs = [q_tok, q_len, a_tok, a_len, targets, ans_tok, ans_len, self.batch, self.step_ctr, self.tokens, self.lens,
                self.frames, self.row_valid, self.neg_idx]
        self._bb_ptrs = (ctypes.c_ulonglong * 18)(*[t.data_ptr() for t in ptrs], 0, 0, 0, 0)
        self._bb_dims = (ctypes.c_int * 8)(targets.shape[1], ans_tok.shape[0], self.L, self.kw, B, K, self.S, 0)

    def train_fwd_bwd(self) -> None:
        """One training step's forward + backward for the samples listed in ``self.batch`` (device), everything on the
        current stream and CUDA-graph capturable: batch assembly with negative sampling, ONE encoder pass over the
        question, the positive answer and the K candidates of every sample, first-violator selection, and the
        backward pass of the selected triples (bicnn.lua:279-420).  Gradients: ``gpk`` (dense, packed) and
        ``gflat[:n_lookup]`` (embedding)."""
        k, st = self.k, cur_stream
        check(k.mpit_bicnn_build_batch(self._bb_ptrs, self._bb_dims, self.seed, st()), "mpit_bicnn_build_batch")
        self.encode()
        check(k.mpit_gesd_pairs(ptr(self.y), ptr(self.qi_all), ptr(self.ai_all), ptr(self.sim_all), ptr(self.stats_all),
                                self.qi_all.numel(), self.F, self.Fp, st()), "mpit_gesd_pairs")
        check(k.mpit_bicnn_select(ptr(self.sim_all), ptr(self.stats_all), ptr(self.batch), self.B, self.K, self.margin,
                                  ptr(self.ai2), ptr(self.sim2), ptr(self.stats2), ptr(self.active), ptr(self.counters),
                                  ptr(self.step_ctr), st()), "mpit_bicnn_select")
        self.gflat[:self.n_lookup].zero_()
        self.backward_from_pairs(self.qi2, self.ai2, self.sim2, self.stats2, self.gflat, self.loss_acc)

    def _wgrad(self, w):
        k = self.k
        k.mpit_set_tf32x3(0)
        check(k.mpit_umma_wgrad(w["tx"], w["tz"], w["geom"], ptr(self.part), None, ptr(self.err), w["block_n"],
                                w["nsplit"], cur_stream()), "mpit_umma_wgrad")
        check(k.mpit_wgrad_reduce(ptr(self.part), ptr(w["gw"]), w["nsplit"], w["m_pad"], w["chunks"] * 32, w["N"],
                                  w["chunks"] * 32, cur_stream()), "mpit_wgrad_reduce")

    def check(self):
        e = int(self.err.item())
        if e:
            raise RuntimeError(f"umma pipeline wait timed out (code {e})")
