"""CIFAR-10 in the Torch7 layout the reference reads.

``goot.lua:62-69`` does ``torch.load('../cifar-10-batches-t7/data_batch_N.t7', 'ascii')`` for N = 1..5 and
``test_batch.t7``; every file is a table ``{data = ByteTensor(3072, 10000), labels = ByteTensor(1, 10000)}``
(one image per COLUMN -- hence the ``:t()``), labels 0-based in the file (the Lua code adds 1; here they stay
0-based).  The stand-alone loader ``cifar10.lua:40-77`` reads the same files and reshapes to ``[n,3,32,32]``.

The engines take uint8 images and do the ``/255`` normalisation (``goot.lua:80-81``) inside the input kernel, so
the loader returns ``uint8 [n,3,32,32]`` + ``int32 [n]``.  When the files are absent (they are not shipped with
the reference: ``.MISSING_LARGE_BLOBS``) ``load_cifar10`` falls back to synthetic data of the same shape unless
``synthetic_ok=False``.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from .. import t7

CLASSES = ["airplane", "automobile", "bird", "cat", "deer", "dog", "frog", "horse", "ship", "truck"]   # goot.lua:87
_TRAIN_FILES = [f"data_batch_{i}.t7" for i in range(1, 6)]
_TEST_FILE = "test_batch.t7"


def find_cifar10(root: Optional[str] = None) -> Optional[str]:
    """Directory holding the ``*.t7`` batches: ``root``, $CIFAR10_T7, ./cifar-10-batches-t7, ../cifar-10-batches-t7."""
    for d in (root, os.environ.get("CIFAR10_T7"), "cifar-10-batches-t7", "../cifar-10-batches-t7"):
        if d and os.path.exists(os.path.join(d, _TEST_FILE)):
            return d
    return None


def _read_batch(path: str) -> Tuple[torch.Tensor, torch.Tensor]:
    with open(path, "rb") as f:
        head = f.read(2)
    mode = "ascii" if head[:1].isdigit() and head[1:2] in (b"\n", b" ") else "binary"
    tab = t7.load(path, mode)
    data, labels = tab["data"], tab["labels"]
    if data.dim() == 2 and data.shape[0] == 3072:          # [3072, n]: one image per column (goot.lua:64)
        data = data.t()
    x = data.contiguous().view(-1, 3, 32, 32).to(torch.uint8)
    y = labels.reshape(-1).to(torch.int32)                  # labels[1] in Lua (cifar10.lua:57)
    return x, y


def load_cifar10(root: Optional[str] = None, trsize: int = 50000, tesize: int = 10000, synthetic_ok: bool = True,
                 seed: int = 0):
    """-> ((train_x, train_y), (test_x, test_y), is_real).  ``trsize``/``tesize`` truncate like goot.lua:73-77."""
    d = find_cifar10(root)
    if d is None:
        if not synthetic_ok:
            raise FileNotFoundError("cifar-10-batches-t7 not found (set CIFAR10_T7)")
        g = torch.Generator().manual_seed(seed)
        mk = lambda n: (torch.randint(0, 256, (n, 3, 32, 32), dtype=torch.uint8, generator=g),
                        torch.randint(0, 10, (n,), dtype=torch.int32, generator=g))
        return mk(trsize), mk(tesize), False
    xs, ys = [], []
    need = trsize
    for name in _TRAIN_FILES:
        if need <= 0:
            break
        x, y = _read_batch(os.path.join(d, name))
        xs.append(x[:need])
        ys.append(y[:need])
        need -= xs[-1].shape[0]
    tx, ty = _read_batch(os.path.join(d, _TEST_FILE))
    return (torch.cat(xs), torch.cat(ys)), (tx[:tesize], ty[:tesize]), True


def write_cifar10_t7(root: str, train: Tuple[torch.Tensor, torch.Tensor], test: Tuple[torch.Tensor, torch.Tensor],
                     mode: str = "ascii") -> None:
    """Write uint8 images / labels in the reference's file layout (five training batches + test batch)."""
    os.makedirs(root, exist_ok=True)
    x, y = train
    per = (x.shape[0] + 4) // 5
    for i, name in enumerate(_TRAIN_FILES):
        xb, yb = x[i * per:(i + 1) * per], y[i * per:(i + 1) * per]
        t7.save(os.path.join(root, name), {"data": xb.reshape(xb.shape[0], -1).t().contiguous().to(torch.uint8),
                                           "labels": yb.reshape(1, -1).to(torch.uint8)}, mode)
    xb, yb = test
    t7.save(os.path.join(root, _TEST_FILE), {"data": xb.reshape(xb.shape[0], -1).t().contiguous().to(torch.uint8),
                                             "labels": yb.reshape(1, -1).to(torch.uint8)}, mode)
