"""Centre variable sharded over the GPUs of one box (device path of the parameter server).

Reference: a set of CPU server processes each owning a contiguous slice of the flat parameter
vector (asyncsgd/pserver.lua:38-62, partition in asyncsgd/pclient.lua:116-129).  Here every GPU is
a worker *and* the owner of shard ``rank``; clients reach every shard with one-sided NVLink
accesses issued by the fused kernels in ``csrc/kernels/comm.cu``.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from .._native import kernels, check, ptr, cur_stream
from .symm import SymmBuffer


def partition(n: int, nshards: int) -> List[Tuple[int, int]]:
    """Reference partition, 0-based ``(offset, size)`` per shard (pclient.lua:118-128: equal
    ``floor(n/S)`` pieces, the last one extended to ``n``; the wire format there is 1-based)."""
    base = n // nshards
    out = []
    for s in range(nshards):
        off = s * base
        size = base if s < nshards - 1 else n - off
        out.append((off, size))
    return out


def shard_capacity(n: int, nshards: int) -> int:
    """Floats to allocate per shard: the largest shard, padded to 4 floats (16-byte quads)."""
    cap = max(sz for _, sz in partition(n, nshards))
    return (cap + 3) // 4 * 4 + 4


class ShardedCenter:
    """Handle on the sharded centre: thin launcher of the exchange kernels."""

    def __init__(self, buf: SymmBuffer, n: int):
        self.buf = buf
        self.n = n
        self.nranks = buf.nranks
        self.rank = buf.rank
        self.parts = partition(n, self.nranks)
        self._ptrs = buf.ptr_array()
        self.k = kernels()
        # large, 16-byte-aligned problems go through the bulk-copy engine (comm_bulk.cu)
        self.bulk = bool(self.k.mpit_bulk_supported(n, self.nranks)) and n >= (1 << 20)

    @classmethod
    def create(cls, heap, n: int) -> "ShardedCenter":
        return cls(heap.alloc(shard_capacity(n, heap.nranks)), n)

    def _nmaps(self, wmap) -> int:
        return 0 if wmap is None else max(1, int(wmap.numel()) // self.n)

    @property
    def local_shard(self) -> torch.Tensor:
        off, size = self.parts[self.rank]
        return self.buf.local[:size]

    # -- one-sided operations ------------------------------------------------------------------
    def put(self, src: torch.Tensor) -> None:
        """centre := src (seeding by the first worker, pclient.lua:131-133)."""
        check(self.k.mpit_param_put(self._ptrs, self.nranks, self.rank, ptr(src), self.n, cur_stream()),
              "mpit_param_put")

    def pull(self, dst: torch.Tensor, wpk=None, wmap=None, map_base: int = 0) -> None:
        """dst := centre (pclient.lua:74-84), optionally refreshing the packed operand copy (``wmap`` covers the
        elements from ``map_base`` on)."""
        if self.bulk and wpk is None and dst.data_ptr() % 16 == 0:
            check(self.k.mpit_param_pull_bulk(self._ptrs, self.nranks, self.rank, ptr(dst), self.n, cur_stream()),
                  "mpit_param_pull_bulk")
            return
        nmaps = 0 if wmap is None else max(1, int(wmap.numel()) // (self.n - map_base))
        check(self.k.mpit_param_pull(self._ptrs, self.nranks, self.rank, ptr(dst), self.n, ptr(wpk), ptr(wmap),
                                     nmaps, map_base, cur_stream()), "mpit_param_pull")

    def push_add(self, src: torch.Tensor, params: Optional[torch.Tensor] = None, *, scale=1.0, l1=0.0, l2=0.0,
                 clip=0.0, zero_src=False) -> None:
        """centre += scale * clip(reg(src)) (pserver.lua:91)."""
        if self.bulk and l1 == 0 and l2 == 0 and clip == 0 and src.data_ptr() % 16 == 0:
            check(self.k.mpit_push_add_bulk(self._ptrs, self.nranks, self.rank, ptr(src), self.n, scale,
                                            int(zero_src), cur_stream()), "mpit_push_add_bulk")
            return
        check(self.k.mpit_push_add(self._ptrs, self.nranks, self.rank, ptr(src), ptr(params), self.n, scale, l1,
                                   l2, clip, int(zero_src), cur_stream()), "mpit_push_add")

    def elastic(self, x: torch.Tensor, d: torch.Tensor, alpha: float) -> None:
        """d = alpha (x - centre); centre += d (optim-eamsgd.lua:58-65)."""
        if self.bulk and x.data_ptr() % 16 == 0 and d.data_ptr() % 16 == 0:
            check(self.k.mpit_elastic_exchange_bulk(self._ptrs, self.nranks, self.rank, ptr(x), ptr(d), self.n, alpha,
                                                    cur_stream()), "mpit_elastic_exchange_bulk")
            return
        check(self.k.mpit_elastic_exchange(self._ptrs, self.nranks, self.rank, ptr(x), ptr(d), self.n, alpha,
                                           cur_stream()), "mpit_elastic_exchange")

    def eamsgd_step(self, x, v, g, gmap, d, wpk, wmap, *, alpha, clr, l2wd, mom_next, do_apply, use_mom,
                    do_lookahead, do_sync, has_d_in, la_in=False) -> None:
        check(self.k.mpit_eamsgd_fused_step(self._ptrs, self.nranks, self.rank, ptr(x), ptr(v), ptr(g), ptr(gmap),
                                            ptr(d), ptr(wpk), ptr(wmap), self._nmaps(wmap), self.n, alpha, clr, l2wd, mom_next,
                                            int(do_apply), int(use_mom), int(do_lookahead), int(do_sync),
                                            int(has_d_in), int(la_in), cur_stream()), "mpit_eamsgd_fused_step")

    # -- checkpoint of the centre: every owner writes / reads its own shard (utils/sharded_ckpt.py) ---------------
    def save(self, path: str, barrier=None) -> None:
        """One Torch7 flat-tensor file holding the whole centre, written in parallel by the shard owners (nothing is
        gathered; ``t7.load`` / ``-loadmodel`` read it).  ``barrier``: e.g. ``torch.distributed.barrier`` when the host
        message layer is not initialised.  The reference never saves the centre."""
        from ..utils.sharded_ckpt import save_flat_sharded
        off, size = self.parts[self.rank]
        torch.cuda.synchronize(self.buf.local.device)
        save_flat_sharded(path, self.buf.local[:size].contiguous(), off, self.n, barrier=barrier, rank=self.rank)

    def load(self, path: str, barrier=None) -> None:
        """Restore this rank's shard from a checkpoint written by :meth:`save` (or any flat FloatTensor of ``n``
        elements, e.g. a parameter file the reference wrote)."""
        from ..utils.sharded_ckpt import load_flat_sharded
        off, size = self.parts[self.rank]
        piece = torch.empty(size, dtype=torch.float32)
        total = load_flat_sharded(path, piece, off, collective=False)
        if total != self.n:
            raise ValueError(f"{path} holds {total} elements, the centre has {self.n}")
        self.buf.local[:size].copy_(piece)
        torch.cuda.synchronize(self.buf.local.device)
        if barrier is not None:
            barrier()

    def gather_full(self) -> torch.Tensor:
        """Materialise the whole centre on this rank (tester role, BiCNN/bicnn.lua:580-596)."""
        out = torch.empty(self.n, dtype=torch.float32, device=self.buf.local.device)
        self.pull(out)
        return out
