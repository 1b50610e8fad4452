"""Sharded checkpoints: every rank writes its slice of a flat fp32 vector into ONE Torch7 file, in parallel.

The reference checkpoints by gathering: the tester rank pulls the whole parameter vector from the servers and
``torch.save``s it (/root/reference/BiCNN/bicnn.lua:580-596); the centre variable and the server-side optimizer
state, which live sharded on the servers, are never saved.  Here the owners write their shards
themselves through the MPI-IO layer (``mpit_b200.msg.fileio``: one collective open, a view that starts after the
Torch7 header, ``File_write_at_all`` at the shard's offset), and the result is byte-for-byte the file ``torch.save`` of
the gathered tensor would be -- so ``-loadmodel`` / ``load_checkpoint`` / ``t7.load`` read it unchanged.  Device
shards are staged by the datatype engine; nothing is gathered on one rank.

Two ways to coordinate: with the host message layer initialised the file routines are collective over ``comm``;
without it (GPU launchers whose control plane is ``torch.distributed``) pass ``barrier=dist.barrier`` and every rank
opens the file on its own.
"""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np
import torch

from .. import t7


def save_flat_sharded(path: str, shard: torch.Tensor, offset: int, total: int, comm: Optional[int] = None,
                      barrier: Optional[Callable[[], None]] = None, rank: Optional[int] = None) -> None:
    """``shard`` = elements ``[offset, offset + shard.numel())`` of a flat fp32 vector of ``total`` elements."""
    from .. import msg as mpiT
    assert shard.dtype == torch.float32 and shard.is_contiguous()
    header = np.frombuffer(t7.flat_float_header(total), dtype=np.uint8).copy()
    nbytes = header.size + 4 * total
    if mpiT.Initialized() and barrier is None:
        # collective over the message layer's communicator
        comm = mpiT.COMM_WORLD if comm is None else comm
        fh = mpiT.File_open(comm, path, mpiT.MODE_CREATE | mpiT.MODE_WRONLY)
        mpiT.File_set_size(fh, nbytes)                    # also cuts off a longer previous checkpoint
        if mpiT.Comm_rank(comm) == 0:
            mpiT.File_write_at(fh, 0, header, header.size, mpiT.BYTE)
        mpiT.File_set_view(fh, header.size, mpiT.FLOAT, mpiT.FLOAT)
        mpiT.File_write_at_all(fh, offset, shard, shard.numel(), mpiT.FLOAT)
        mpiT.File_close(fh)
        return
    # every rank on its own, ordered by the caller's barrier (a single process needs none)
    me = 0 if rank is None else rank
    sync = barrier if barrier is not None else (lambda: None)
    if me != 0 and barrier is None:
        raise ValueError("save_flat_sharded: without the message layer, ranks other than 0 need a barrier")
    if me == 0:
        fh = mpiT.File_open(mpiT.COMM_SELF, path, mpiT.MODE_CREATE | mpiT.MODE_WRONLY)
        mpiT.File_set_size(fh, nbytes)
        mpiT.File_write_at(fh, 0, header, header.size, mpiT.BYTE)
        sync()                                            # the file exists at its final size
    else:
        sync()
        fh = mpiT.File_open(mpiT.COMM_SELF, path, mpiT.MODE_WRONLY)
    mpiT.File_set_view(fh, header.size, mpiT.FLOAT, mpiT.FLOAT)
    mpiT.File_write_at(fh, offset, shard, shard.numel(), mpiT.FLOAT)
    mpiT.File_close(fh)
    sync()                                                # everybody's slice is in the file


def load_flat_sharded(path: str, out: torch.Tensor, offset: int, comm: Optional[int] = None,
                      collective: Optional[bool] = None) -> int:
    """Read elements ``[offset, offset + out.numel())`` of the checkpoint's vector into ``out``; returns the length of
    the whole vector.  Collective over ``comm`` when the message layer is initialised (or ``collective=True``)."""
    from .. import msg as mpiT
    assert out.dtype == torch.float32 and out.is_contiguous()
    if collective is None:
        collective = mpiT.Initialized()
    comm = (mpiT.COMM_WORLD if comm is None else comm) if collective else mpiT.COMM_SELF
    fh = mpiT.File_open(comm, path, mpiT.MODE_RDONLY)
    size = mpiT.File_get_size(fh)
    # the header length depends on nothing but the element count: find the count that explains the file size
    total = (size - len(t7.flat_float_header(0))) // 4
    header = t7.flat_float_header(total)
    probe = np.zeros(len(header), dtype=np.uint8)
    mpiT.File_read_at(fh, 0, probe, probe.size, mpiT.BYTE)
    if probe.tobytes() != header or len(header) + 4 * total != size:
        mpiT.File_close(fh)
        raise ValueError(f"{path} is not a flat torch.FloatTensor checkpoint")
    if offset + out.numel() > total:
        mpiT.File_close(fh)
        raise ValueError(f"slice [{offset}, {offset + out.numel()}) exceeds the checkpoint's {total} elements")
    mpiT.File_set_view(fh, len(header), mpiT.FLOAT, mpiT.FLOAT)
    mpiT.File_read_at_all(fh, offset, out, out.numel(), mpiT.FLOAT)
    mpiT.File_close(fh)
    return total
