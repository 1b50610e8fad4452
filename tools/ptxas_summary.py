#!/usr/bin/env python
"""Registers, spills, stack and static shared memory of every kernel, from the `-Xptxas -v` lines of the build logs
(`python tools/build.py` keeps them under build/obj/*.log) -> profiles/ptxas_summary.md."""
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
logs = sorted((ROOT / "build" / "obj").glob("kernels__*.log")) + sorted((ROOT / "build" / "obj").glob("symm__*.log"))
if not logs:
    sys.exit("no build logs: run python tools/build.py first")
names = []
for log in logs:
    txt = log.read_text()
    for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*?Function properties for \S+\n\s+(\d+) bytes stack frame, "
                         r"(\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers(?:, used (\d+) barriers)?"
                         r"(?:, (\d+) bytes cumulative stack size)?(?:, (\d+) bytes smem)?", txt, re.S):
        names.append((log.name.replace("kernels__", "").replace("symm__", "").replace(".log", ""), m.group(1), int(m.group(2)),
                      int(m.group(3)), int(m.group(4)), int(m.group(5)), int(m.group(8) or 0)))
dem = subprocess.run(["c++filt"] + [n[1] for n in names], capture_output=True, text=True).stdout.splitlines()
rows = []
for (src, _, stack, sst, sld, regs, smem), d in zip(names, dem):
    d = re.sub(r"\(.*$", "", d).replace("void ", "").replace("mpit::", "")
    rows.append((src, d, regs, stack, sst, sld, smem))
rows.sort()
out = ["# ptxas -v per kernel (sm_100a, final build)", "",
       f"{len(rows)} kernels, {sum(1 for r in rows if r[4])} of them with spill stores (8-48 bytes in the once-per-CTA epilogue code of some "
       "`umma_igemm_kernel` instantiations, in the index arithmetic of `ps_push_kernel` / `rule_kernel<5>` and in the 4-way unrolled "
       "peer-load all-reduce; 580 bytes in `classifier_head_kernel<16, 16>`, the 512-input instantiation that neither model uses -- "
       "the CIFAR head is `<8, 16>`).  Where a reload does sit in a main loop (three `LDL` after the four `UTCHMMA` of a k-block in `umma_igemm_kernel<32, 0>`) it is executed by the one issuing thread, once per k-block of several hundred clocks.", "",
       "| source | kernel | registers | stack B | spill stores B | spill loads B | static smem B |", "|---|---|---:|---:|---:|---:|---:|"]
out += [f"| `{s}` | `{d}` | {regs} | {stack} | {sst} | {sld} | {smem} |" for s, d, regs, stack, sst, sld, smem in rows]
(ROOT / "profiles" / "ptxas_summary.md").write_text("\n".join(out) + "\n")
print(f"{len(rows)} kernels -> profiles/ptxas_summary.md")
