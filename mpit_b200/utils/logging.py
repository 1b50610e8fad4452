"""Headless stand-ins for ``optim.Logger``, ``optim.ConfusionMatrix``, ``xlua.progress`` and the
``tm.*`` timing buckets of the reference loops.  Plotting (gnuplot, goot.lua:239-244) is dropped:
logs go to files."""
from __future__ import annotations

import os
import sys
import time
from contextlib import contextmanager
from typing import Dict, List, Optional

import torch


class Logger:
    """Tab-separated column log, one ``add`` per row (optim.Logger; goot.lua:96-97,236-238)."""

    def __init__(self, path: str):
        self.path = path
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        self.names: Optional[List[str]] = None
        self.f = open(path, "w")

    def add(self, row: Dict[str, object]) -> None:
        if self.names is None:
            self.names = list(row)
            self.f.write("\t".join(self.names) + "\n")
        self.f.write("\t".join(str(row[k]) for k in self.names) + "\n")
        self.f.flush()

    def style(self, *_a, **_k):   # accepted and ignored: no display on this box
        pass

    def plot(self):
        pass

    def close(self):
        self.f.close()


class ConfusionMatrix:
    """optim.ConfusionMatrix (goot.lua:93,209,232): ``add(output, target)``, ``totalValid``, ``zero``."""

    def __init__(self, classes):
        self.classes = list(classes)
        n = len(self.classes)
        self.mat = torch.zeros(n, n, dtype=torch.int64)

    def add(self, output: torch.Tensor, target: int) -> None:
        self.mat[int(target), int(output.argmax())] += 1

    def batch_add(self, outputs: torch.Tensor, targets: torch.Tensor) -> None:
        pred = outputs.argmax(dim=1).cpu().long()
        idx = targets.cpu().long() * len(self.classes) + pred
        self.mat += torch.bincount(idx, minlength=self.mat.numel()).view_as(self.mat)

    def merge(self, counts: torch.Tensor) -> None:
        self.mat += counts.cpu().long()

    @property
    def totalValid(self) -> float:
        tot = int(self.mat.sum())
        return float(self.mat.diag().sum()) / tot if tot else 0.0

    def zero(self) -> None:
        self.mat.zero_()

    def __str__(self):
        rows = ["ConfusionMatrix:"]
        for i, c in enumerate(self.classes):
            r = self.mat[i]
            acc = float(r[i]) / max(int(r.sum()), 1) * 100
            rows.append("[" + " ".join(f"{int(v):7d}" for v in r) + f"]  {acc:6.2f}%  [class: {c}]")
        rows.append(f" + global correct: {self.totalValid * 100:.4f}%")
        return "\n".join(rows)


def _nvtx():
    """torch.cuda.nvtx when a CUDA runtime is present and tracing is wanted (MPIT_NVTX=1), else None."""
    if os.environ.get("MPIT_NVTX", "0") != "1":
        return None
    try:
        import torch
        return torch.cuda.nvtx if torch.cuda.is_available() else None
    except Exception:       # pragma: no cover
        return None


class Timers(dict):
    """Accumulating wall-clock buckets (``tm.feval``, ``tm.sync``, ...; goot.lua:24-26, bicnn.lua:17-28).

    With ``MPIT_NVTX=1`` every bucket is also an NVTX range, so the same compute / sync / transfer split shows up
    on an Nsight timeline.  ``device=True`` buckets are timed with CUDA events on the current stream instead of the
    host clock (resolved lazily by :meth:`resolve`, so that timing never synchronises the step)."""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self._pending = []

    @contextmanager
    def __call__(self, name: str, device: bool = False):
        nv = _nvtx()
        if nv is not None:
            nv.range_push(name)
        if device:
            import torch
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        t0 = time.perf_counter()
        try:
            yield
        finally:
            if device:
                e1.record()
                self._pending.append((name, e0, e1))
            else:
                self[name] = self.get(name, 0.0) + time.perf_counter() - t0
            if nv is not None:
                nv.range_pop()

    def resolve(self) -> "Timers":
        """Fold finished device-timed ranges into their buckets (seconds).  Synchronises on their end events."""
        for name, e0, e1 in self._pending:
            e1.synchronize()
            self[name] = self.get(name, 0.0) + e0.elapsed_time(e1) / 1e3
        self._pending = []
        return self


def progress(i: int, n: int, every: int = 0, stream=sys.stderr) -> None:
    """xlua.progress (goot.lua:169,279), as a sparse one-line counter."""
    if every and (i % every == 0 or i == n):
        stream.write(f"\r[{i}/{n}]")
        stream.flush()
        if i == n:
            stream.write("\n")
