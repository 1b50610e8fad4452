"""Linear layer whose weights are pulled from the sharded centre *inside* the GEMM (north-star item:
"the parameter pull, fused with the first GEMM that consumes the fetched weights").

The weight matrix ``W [out, in]`` is row-sharded over the ranks' symmetric memory (rank s owns rows
``[s*out/S, (s+1)*out/S)``, i.e. the reference's contiguous partition of the flat Linear weight when
``out % S == 0``).  ``forward`` launches ONE tcgen05 kernel whose B-operand TMA loads read every weight
tile from its owner's shard over NVLink while the tensor cores work on earlier tiles; no copy of W is
ever materialised locally.  The un-fused formulation is ``ShardedCenter.pull`` followed by a GEMM.

Measured (tools/bench_pulled_linear.py): 1.5-3.4x faster than NCCL all_gather + cuBLAS at 2 and 8
GPUs while the batch is one or two 128-row tiles; every further M-tile fetches its weight tiles again, so for
M >= 512 the un-fused form wins.  The remedy is reuse inside the CTA (one CTA owning several 128-row M-tiles, i.e.
more TMEM accumulator columns per fetched weight tile) or bringing the shard into local memory first; TMA multicast
loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.
"""
from __future__ import annotations

import ctypes

import torch

from .._native import kernels, check, ptr, cur_stream
from .umma_ops import make_tmap, igemm_geom, weight_tmap, EPI_STORE


class PulledLinear:
    def __init__(self, heap, in_features: int, out_features: int, max_rows: int, act: int = 0):
        S = heap.nranks
        assert in_features % 32 == 0 and out_features % (32 * S) == 0
        self.S, self.K, self.N = S, in_features, out_features
        self.rows_per_owner = out_features // S
        self.block_n = 64 if self.rows_per_owner % 64 == 0 else 32
        self.shard = heap.alloc(self.rows_per_owner * in_features)       # my rows of W (symmetric)
        self.M = (max_rows + 127) // 128 * 128
        dev = self.shard.local.device
        self.x = torch.zeros(self.M * in_features, device=dev)
        self.y = torch.zeros(self.M * out_features, device=dev)
        self.err = torch.zeros(1, dtype=torch.int32, device=dev)
        self.act = act
        maps = (ctypes.c_ubyte * (128 * S))()
        for r, p in enumerate(self.shard.ptrs):       # one tensor map per owner, over the peer-mapped shard
            m = make_tmap(p, [in_features, self.rows_per_owner], [in_features * 4], [32, self.block_n])
            ctypes.memmove(ctypes.addressof(maps) + 128 * r, m, 128)
        self._maps = maps
        self._ta = make_tmap(self.x.data_ptr(), [in_features, 1, 1, self.M], [in_features * 4] * 3, [32, 1, 1, 128])
        self._geom = igemm_geom(nsub=1, tw=1, th=1, tn=128, tiles_w=1, tiles_h=1, kw_it=1, cb=in_features // 32,
                                num_kb=in_features // 32, sgn=1, B=self.M, OH=1, OW=1, N=out_features,
                                mode=EPI_STORE, relu=act)
        # Un-fused fallback for batches of more than FUSED_MAX_TILES 128-row tiles (every M-tile of the fused kernel
        # fetches its weight tiles over NVLink again): gather W once with in-kernel peer loads, then a local GEMM.
        self._wfull = None
        self._tb_local = None

    FUSED_MAX_TILES = 2

    def _forward_unfused(self, bias):
        k = kernels()
        if self._wfull is None:
            self._wfull = torch.empty(self.N * self.K, device=self.x.device)
            self._tb_local = weight_tmap(self._wfull, self.N, self.K, self.block_n)
        n = self.rows_per_owner * self.K
        for r, p in enumerate(self.shard.ptrs):                       # MPI_Get of every owner's rows (ld.sys over NVLink)
            check(k.mpit_peer_rma(1, p, self._wfull.data_ptr() + 4 * r * n, n, 1.0, cur_stream()), "mpit_peer_rma")
        check(k.mpit_umma_igemm(self._ta, self._tb_local, self._geom, 1.0, ptr(bias), ptr(self.y), None, None, None, None,
                                None, ptr(self.err), None, self.block_n, cur_stream()), "mpit_umma_igemm")

    def weight_rows(self) -> torch.Tensor:
        """This rank's rows of W, as a [rows_per_owner, in] view of its symmetric shard."""
        return self.shard.local[:self.rows_per_owner * self.K].view(self.rows_per_owner, self.K)

    def forward(self, x: torch.Tensor, bias: torch.Tensor = None, fused: bool = None) -> torch.Tensor:
        """``fused=None``: the fused kernel for up to FUSED_MAX_TILES row tiles, gather + GEMM beyond."""
        m = x.shape[0]
        self.x[:m * self.K].copy_(x.reshape(-1))
        if fused is None:
            fused = (m + 127) // 128 <= self.FUSED_MAX_TILES
        if not fused:
            self._forward_unfused(bias)
            return self.y.view(self.M, self.N)[:m]
        check(kernels().mpit_umma_linear_pull(self._ta, self._maps, self.S, self.rows_per_owner, self._geom, ptr(bias),
                                              ptr(self.y), ptr(self.err), self.block_n, cur_stream()),
              "mpit_umma_linear_pull")
        return self.y.view(self.M, self.N)[:m]
