#!/usr/bin/env python
"""Which of the reference's MPI wrappers does ``mpit_b200.msg`` provide?

The reference generates one Lua wrapper per JSON file under ``/root/reference/specs`` (270 files,
/root/reference/readspec.py:73-188; it skips a handful whose signatures it cannot express).  This tool reads the
``name`` / ``category`` of every spec and looks the function up on ``mpit_b200.msg``; ``--markdown`` prints the
per-chapter table kept in ``docs/MPI_COVERAGE.md``.
"""
from __future__ import annotations

import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def coverage(spec_dir: str = "/root/reference/specs"):
    import mpit_b200.msg as m
    rows = {}
    for f in sorted(glob.glob(os.path.join(spec_dir, "*.json"))):
        d = json.load(open(f))
        name = d["name"].replace("MPI_", "", 1)
        chapter = d.get("category", "?").replace(" C Bindings", "")
        have = callable(getattr(m, name, None))
        rows.setdefault(chapter, []).append((name, have))
    return rows


def main():
    rows = coverage()
    total = sum(len(v) for v in rows.values())
    have = sum(h for v in rows.values() for _, h in v)
    if "--markdown" in sys.argv:
        print("| chapter | provided / in the reference's specs | missing |")
        print("|---|---:|---|")
        for ch, v in sorted(rows.items()):
            miss = ", ".join(n for n, h in v if not h) or "–"
            print(f"| {ch} | {sum(h for _, h in v)} / {len(v)} | {miss} |")
        print(f"| **total** | **{have} / {total}** | |")
    else:
        for ch, v in sorted(rows.items()):
            miss = [n for n, h in v if not h]
            print(f"{ch}: {sum(h for _, h in v)}/{len(v)}" + (f"  missing: {miss}" if miss else ""))
        print(f"total: {have}/{total}")
    return 0 if have == total else 1


if __name__ == "__main__":
    sys.exit(main())
