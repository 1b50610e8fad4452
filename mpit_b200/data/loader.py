"""Prefetching mini-batch loader on a native worker thread (csrc/io/loader.cpp).

The reference builds each mini-batch inside the training loop, sample by sample (per-epoch ``torch.randperm``,
/root/reference/asyncsgd/goot.lua:164; a random crop position and a copy per sample, goot.lua:172-184).  Here a C++
thread does that ahead of time into a ring of pinned batch slots; :meth:`BatchLoader.next` hands out the next filled
slot as three tensors that alias pinned memory, ready for ``CifarTrainer.train_step``.
"""
from __future__ import annotations

import ctypes
from typing import Tuple

import torch

from .._native import host


class BatchLoader:
    """``images``: uint8 ``[n, ...]`` (contiguous, host), ``labels``: int32 ``[n]``.  ``crop_range`` = number of crop
    positions per axis (32 - 28 + 1 = 5 for CIFAR; 0 or 1 = always zero, i.e. no random crop).

    ``next()`` -> ``(images [B, ...] uint8, labels [B] int32, crop [B, 2] int32)``, views into the ring: they stay
    valid until the following ``next()`` call, which recycles the slot (``train_step`` has copied a batch to the
    device by the time it returns, so calling ``next`` once per step is safe)."""

    def __init__(self, images: torch.Tensor, labels: torch.Tensor, batch: int, *, nslots: int = 4,
                 crop_range: int = 5, seed: int = 0, shuffle: bool = True, pin: bool = True):
        assert images.dtype == torch.uint8 and images.is_contiguous() and not images.is_cuda
        assert labels.dtype == torch.int32 and labels.is_contiguous() and labels.numel() == images.shape[0]
        self.images, self.labels, self.batch, self.nslots = images, labels, batch, nslots
        sample_shape = tuple(images.shape[1:])
        sample_bytes = int(images[0].numel())
        pin = pin and torch.cuda.is_available()
        mk = (lambda *s, dtype: torch.empty(*s, dtype=dtype).pin_memory()) if pin else \
            (lambda *s, dtype: torch.empty(*s, dtype=dtype))
        self.ring_x = mk(nslots, batch, *sample_shape, dtype=torch.uint8)
        self.ring_y = mk(nslots, batch, dtype=torch.int32)
        self.ring_c = mk(nslots, batch, 2, dtype=torch.int32)
        self._h = ctypes.c_void_p(0)
        rc = host().mpit_loader_create(ctypes.c_void_p(images.data_ptr()), ctypes.c_void_p(labels.data_ptr()),
                                       images.shape[0], sample_bytes, batch, nslots,
                                       ctypes.c_void_p(self.ring_x.data_ptr()), ctypes.c_void_p(self.ring_y.data_ptr()),
                                       ctypes.c_void_p(self.ring_c.data_ptr()), crop_range, seed & (2 ** 64 - 1),
                                       int(shuffle), ctypes.byref(self._h))
        if rc != 0:
            raise ValueError(f"mpit_loader_create failed with code {rc}")
        self._held = -1
        self.epoch = 0
        self.batches_per_epoch = images.shape[0] // batch

    def next(self) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        lib = host()
        if self._held >= 0:
            lib.mpit_loader_release(self._h, self._held)
        slot, epoch = ctypes.c_int(-1), ctypes.c_longlong(0)
        rc = lib.mpit_loader_next(self._h, ctypes.byref(slot), ctypes.byref(epoch))
        if rc != 0:
            raise RuntimeError(f"mpit_loader_next failed with code {rc}")
        self._held, self.epoch = slot.value, epoch.value
        s = slot.value
        return self.ring_x[s], self.ring_y[s], self.ring_c[s]

    def __iter__(self):
        while True:
            yield self.next()

    def close(self) -> None:
        if self._h:
            host().mpit_loader_destroy(self._h)
            self._h = ctypes.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
