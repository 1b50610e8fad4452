"""Linear layer whose weights are pulled from the sharded centre *inside* the GEMM (north-star item:
"the parameter pull, fused with the first GEMM that consumes the fetched weights"), with its backward.

The weight matrix ``W [out, in]`` is row-sharded over the ranks' symmetric memory (rank s owns rows
``[s*out/S, (s+1)*out/S)``, i.e. the reference's contiguous partition of the flat Linear weight when
``out % S == 0``; reference sequence: pull, ``/root/reference/asyncsgd/pclient.lua:74-84``, then the first
``nn.Linear`` forward of the model, ``asyncsgd/models/7-layers-cnn.lua:20``).

* ``forward`` launches ONE tcgen05 kernel whose B-operand TMA loads read every weight tile from its owner's shard
  over NVLink while the tensor cores work on earlier tiles; no copy of W is ever materialised locally.  One CTA owns up
  to four 128-row M-tiles of the batch (four TMEM accumulators), so a weight tile crosses NVLink once per CTA however
  large the batch is.  (TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from
  another GPU's memory hung the GPU -- hence reuse across accumulators instead of sharing a tile between CTAs.)
* ``backward`` computes ``dx = dy W`` the same way: the reduction now runs over W's rows, i.e. across owners, and each
  owner keeps its rows a second time transposed (``[in][rows]``) so that the tiles are K-major for the tensor core;
  the owner of a k-block is chosen per k-block.  The output has few tiles and the reduction is long, so the reduction is
  split over up to 8 CTAs per output tile that add into ``dx`` (``red.add.v4.f32``).
* Weight tiles travel in their own ring (up to 16 in flight per CTA), ahead of the activation ring; activations are read
  in place through a tensor map over the caller's tensor (no staging copy, no padding: out-of-bounds rows read as zeros).
* ``weight_grad`` is the local tcgen05 weight-gradient kernel; ``push`` adds ``scale * dW`` into the owners' rows AND
  their transposed copies with ``red.add`` over NVLink (``pclient.lua:49-59`` + ``pserver.lua:91``).

Measured (tools/bench_pulled_linear.py) against NCCL all_gather + cuBLAS: 1.2-3.4x forward, 1.0-2.7x
backward at 8 GPUs.
"""
from __future__ import annotations

import ctypes

import torch

from .._native import kernels, check, ptr, cur_stream
from .umma_ops import make_tmap, igemm_geom, weight_tmap, EPI_STORE

MN_SWIZZLE = 2


def _tiles_per_cta(m_tiles: int, col_tiles: int) -> int:
    """M-tiles per CTA: as many as the batch has (up to four) -- every extra tile is a re-use of the weight tiles that
    crossed NVLink -- but not so many that the grid falls under ~64 CTAs on 148 SMs."""
    for mt in (4, 2):
        if m_tiles >= mt and -(-m_tiles // mt) * col_tiles >= 64:
            return mt
    return 1


def _split_k(ctas: int, num_kb: int) -> int:
    """Split-K factor of the data gradient (few output tiles, long reduction): double while the grid still fits the
    148 SMs and every slice keeps at least eight k-blocks."""
    ks = 1
    while ks < 8 and ctas * ks * 2 <= 148 and num_kb % (ks * 2) == 0 and num_kb // (ks * 2) >= 8:
        ks *= 2
    return ks


class PulledLinear:
    def __init__(self, heap, in_features: int, out_features: int, max_rows: int, act: int = 0):
        S = heap.nranks
        assert in_features % 32 == 0 and out_features % (32 * S) == 0
        self.S, self.K, self.N = S, in_features, out_features
        self.rank = heap.rank
        self.rows_per_owner = rpo = out_features // S
        self.block_n = 64 if rpo % 64 == 0 else 32
        # output-column tile of the data gradient: 64 columns halve the re-reads of dy (split-K supplies the CTAs)
        self.block_k = 64 if in_features % 64 == 0 else 32
        self.shard = heap.alloc(rpo * in_features)                  # my rows of W              [rpo][in]
        self.shard_t = heap.alloc(rpo * in_features)                # the same rows, transposed [in][rpo]
        self.M = (max_rows + 127) // 128 * 128
        dev = self.shard.local.device
        f32 = dict(dtype=torch.float32, device=dev)
        self.x = torch.zeros(self.M * in_features, **f32)
        self.y = torch.zeros(self.M * out_features, **f32)
        self.dy = torch.zeros(self.M * out_features, **f32)
        self.dx = torch.zeros(self.M * in_features, **f32)
        self.gw = torch.zeros(out_features * in_features, **f32)
        self.err = torch.zeros(1, dtype=torch.int32, device=dev)
        self.act = act
        K, N, M = in_features, out_features, self.M

        def owner_maps(buf, inner, rows, box_rows):
            maps = (ctypes.c_ubyte * (128 * S))()
            for r, p in enumerate(buf.ptrs):          # one tensor map per owner, over the peer-mapped shard
                m = make_tmap(p, [inner, rows], [inner * 4], [32, box_rows])
                ctypes.memmove(ctypes.addressof(maps) + 128 * r, m, 128)
            return maps

        self._maps = owner_maps(self.shard, K, rpo, self.block_n)
        self._maps_t = owner_maps(self.shard_t, rpo, K, self.block_k)
        self._tmaps = {}                                             # tensor maps over activations, by (pointer, shape)
        self._x = self._dy = None
        self._geom = igemm_geom(nsub=1, tw=1, th=1, tn=128, tiles_w=1, tiles_h=1, kw_it=1, cb=K // 32,
                                num_kb=K // 32, sgn=1, B=M, OH=1, OW=1, N=N, mode=EPI_STORE, relu=act)
        # weight gradient dW[n][k] = sum_m dy[m][n] x[m][k] on the tensor cores (MN-major operands)
        self._wg = dict(block_n=64 if N % 64 == 0 else 32, m_pad=-(-(K // 32) // 4) * 128)
        self._part = torch.empty(min(8, M // 64) * self._wg["m_pad"] * N, **f32)
        # un-fused fallback (gather with in-kernel peer loads, then a local GEMM), kept for comparison
        self._wfull = None
        self._tb_local = None

    # -- weights -------------------------------------------------------------------------------------------------
    def weight_rows(self) -> torch.Tensor:
        """This rank's rows of W, as a [rows_per_owner, in] view of its symmetric shard."""
        return self.shard.local[:self.rows_per_owner * self.K].view(self.rows_per_owner, self.K)

    def weight_rows_t(self) -> torch.Tensor:
        return self.shard_t.local[:self.rows_per_owner * self.K].view(self.K, self.rows_per_owner)

    def load_rows(self, rows: torch.Tensor) -> None:
        """Set this rank's rows (both layouts)."""
        self.weight_rows().copy_(rows)
        self.weight_rows_t().copy_(rows.t())

    # -- forward ---------------------------------------------------------------------------------------------------
    def _forward_unfused(self, ta, bias):
        k = kernels()
        if self._wfull is None:
            self._wfull = torch.empty(self.N * self.K, device=self.x.device)
            self._tb_local = weight_tmap(self._wfull, self.N, self.K, self.block_n)
        n = self.rows_per_owner * self.K
        for r, p in enumerate(self.shard.ptrs):                       # MPI_Get of every owner's rows (ld.sys over NVLink)
            check(k.mpit_peer_rma(1, p, self._wfull.data_ptr() + 4 * r * n, n, 1.0, cur_stream()), "mpit_peer_rma")
        k.mpit_set_tf32x3(0)
        g = self._geom
        g[10] = self.m                                                # B: rows of this batch
        check(k.mpit_umma_igemm(ta, self._tb_local, g, 1.0, ptr(bias), ptr(self.y), None, None, None, None,
                                None, ptr(self.err), None, self.block_n, cur_stream()), "mpit_umma_igemm")

    # -- operands in place ---------------------------------------------------------------------------------------------
    def _operand(self, t: torch.Tensor, inner: int, stage: torch.Tensor, box_rows: int = 128, swizzle: int = 1):
        """Tensor map over the caller's activations themselves when they can be read in place (fp32, contiguous, on this
        device, 16-byte aligned): rows past the batch are out of bounds for the map and arrive as zeros, so neither a
        staging copy nor padding is needed.  Anything else is copied into the layer's staging buffer first."""
        m = t.shape[0]
        if not (t.dtype == torch.float32 and t.is_contiguous() and t.device == stage.device and t.data_ptr() % 16 == 0):
            stage[:m * inner].copy_(t.reshape(-1))
            t = stage[:m * inner].view(m, inner)
        key = (t.data_ptr(), m, inner, box_rows, swizzle)
        tm = self._tmaps.get(key)
        if tm is None:
            if len(self._tmaps) > 256:
                self._tmaps.clear()
            tm = make_tmap(t.data_ptr(), [inner, 1, 1, m], [inner * 4] * 3, [32, 1, 1, box_rows], swizzle)
            self._tmaps[key] = tm
        return t, tm

    def forward(self, x: torch.Tensor, bias: torch.Tensor = None, fused: bool = True, mt: int = 0) -> torch.Tensor:
        """``y = act(x W^T + b)``.  ``fused=False`` runs the gather + GEMM formulation; ``mt`` overrides the number of
        M-tiles per CTA (default: as many as the batch has, up to four)."""
        m = x.shape[0]
        assert m <= self.M
        self.m = m
        self._x, ta = self._operand(x, self.K, self.x)          # kept for weight_grad()
        if not fused:
            self._forward_unfused(ta, bias)
            return self.y.view(self.M, self.N)[:m]
        mt = mt or _tiles_per_cta((m + 127) // 128, self.N // self.block_n)
        check(kernels().mpit_umma_pulled_gemm(ta, self._maps, self.S, self.rows_per_owner, 0, m, self.K, self.N,
                                              self.act, ptr(bias), ptr(self.y), ptr(self.err), self.block_n, mt, 1,
                                              cur_stream()), "mpit_umma_pulled_gemm")
        return self.y.view(self.M, self.N)[:m]

    # -- backward --------------------------------------------------------------------------------------------------
    def backward(self, dy: torch.Tensor, mt: int = 0, ksplit: int = 0) -> torch.Tensor:
        """``dx = dy W`` for the batch of the last forward; ``dy`` is the gradient w.r.t. the pre-activation (the caller
        applies the activation's derivative).  W's transposed shards are read from their owners inside the GEMM.  The
        output has few tiles and the reduction is long (all of W's rows), so all M-tiles (up to four) share one CTA --
        W crosses NVLink once -- and the reduction is cut over ``ksplit`` CTAs that add into ``dx``."""
        m = dy.shape[0]
        assert m == self.m
        self._dy, ta = self._operand(dy, self.N, self.dy)
        m_tiles = (m + 127) // 128
        mt = mt or (4 if m_tiles > 2 else m_tiles)
        ksplit = ksplit or _split_k(-(-m_tiles // mt) * (self.K // self.block_k), self.N // 32)
        check(kernels().mpit_umma_pulled_gemm(ta, self._maps_t, self.S, self.rows_per_owner, 1, m, self.N,
                                              self.K, 0, None, ptr(self.dx), ptr(self.err), self.block_k, mt, ksplit,
                                              cur_stream()), "mpit_umma_pulled_gemm")
        return self.dx.view(self.M, self.K)[:m]

    def weight_grad(self) -> torch.Tensor:
        """``dW [out, in]`` of the last forward / backward pair (local: this rank's batch)."""
        k, w = kernels(), self._wg
        _, tx = self._operand(self._x, self.K, self.x, 64, MN_SWIZZLE)
        _, tz = self._operand(self._dy, self.N, self.dy, 64, MN_SWIZZLE)
        nkb = (self.m + 63) // 64
        nsplit = min(8, nkb)
        geom = (ctypes.c_int * 11)(1, 1, 64, 1, 1, 1, self.K // 32, self.K // 32, nkb, self.m, self.N)
        k.mpit_set_tf32x3(0)
        check(k.mpit_umma_wgrad(tx, tz, geom, ptr(self._part), None, ptr(self.err), w["block_n"], nsplit, cur_stream()),
              "mpit_umma_wgrad")
        check(k.mpit_wgrad_reduce(ptr(self._part), ptr(self.gw), nsplit, w["m_pad"], self.K, self.N, self.K,
                                  cur_stream()), "mpit_wgrad_reduce")
        return self.gw.view(self.N, self.K)

    def push(self, scale: float) -> None:
        """Owners' rows (and transposed copies) ``+= scale * dW``: a plain SGD step is ``push(-lr)``; every rank's
        contribution accumulates (``p += g`` on the server, pserver.lua:91)."""
        check(kernels().mpit_pulled_linear_push(self.shard.ptr_array(), self.shard_t.ptr_array(), self.S, self.rank,
                                                ptr(self.gw), self.N, self.K, scale, cur_stream()),
              "mpit_pulled_linear_push")

    def check(self) -> None:
        e = int(self.err.item())
        if e:
            raise RuntimeError(f"umma pipeline wait timed out (code {e})")
