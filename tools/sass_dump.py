#!/usr/bin/env python
"""Dump SASS evidence for every kernel of libmpit_kernels.so: a per-kernel table of the instructions
that prove the Blackwell path (UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA, RED/ATOM,
multimem) into profiles/sass_summary.md, and full listings of the tensor-core kernels under profiles/sass/."""
import os
import re
import subprocess
import sys
from collections import Counter

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "mpit_b200", "lib", "libmpit_kernels.so")
OUT = os.path.join(ROOT, "profiles")
KEYS = [("UTC.MMA", r"UTC[A-Z]*MMA"), ("LDTM", r"\bLDTM"), ("UTMALDG", r"UTMALDG"), ("UTCBAR", r"UTCBAR"),
        ("SYNCS", r"SYNCS"), ("REDG", r"\bREDG"), ("REDG.x4.SYS", r"REDG\.E\.ADD\.F32x4[A-Z.]*SYS"), ("LDGMC", r"\bLDGMC"), ("MULTIMEM", r"LDGMC|STG[A-Z0-9.]*MC|RED[A-Z0-9.]*MC|\.MC\b"),
        ("ATOMG", r"\bATOMG"), ("LDG.SYS", r"LDG[A-Z0-9.]*STRONG\.SYS"), ("STG.SYS", r"STG[A-Z0-9.]*STRONG\.SYS"), ("LDG", r"\bLDG"), ("STG", r"\bSTG"),
        ("FFMA", r"\bFFMA"), ("SHFL", r"\bSHFL"), ("HMMA", r"\bHMMA")]


LISTED = ("umma_igemm_kernel<64, 1, false>", "umma_igemm_kernel<32, 0, false>", "umma_igemm_kernel<32, 2, false>",
          "umma_igemm_kernel<64, 3, false>", "umma_igemm_kernel<64, 1, true>", "umma_igemm_kernel<64, 3, true>",
          "umma_wgrad_kernel", "umma_resident_kernel<64, 3>", "umma_igemm_pull_kernel<32>",
          "umma_pulled_gemm_kernel<64, 4, false>", "umma_pulled_gemm_kernel<32, 1, true>")
# exchange / collective / server kernels whose listings are kept as well (no tensor-core instructions in them)
LISTED_COMM = ("allreduce_twoshot_kernel<true, 8>", "allreduce_oneshot_kernel", "exchange_kernel<4>", "exchange_kernel<5>",
               "ps_push_kernel", "ps_poll_kernel", "ps_apply_kernel<2>", "ps_wait_kernel", "mbx_copy_kernel",
               "mbx_publish_kernel", "pulled_push_kernel", "bicnn_build_batch_kernel", "bicnn_select_kernel",
               "multi_wgrad_reduce_kernel")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = re.split(r"\n\s*Function : ", sass)[1:]
    os.makedirs(os.path.join(OUT, "sass"), exist_ok=True)
    rows = []
    for k in kernels:
        name_m, body = k.split("\n", 1)
        name = subprocess.run(["c++filt", name_m.strip()], capture_output=True, text=True).stdout.strip() or name_m
        short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "")).replace("void ", "").replace("mpit::", "")
        cnt = Counter()
        for label, pat in KEYS:
            cnt[label] = len(re.findall(pat, body))
        ninstr = len(re.findall(r"^\s+/\*[0-9a-f]{4}\*/", body, flags=re.M))
        rows.append((short, ninstr, cnt))
        # full listings (gzip) only for the tensor-core instantiations the CIFAR step actually launches
        if (cnt["UTC.MMA"] and any(t in short for t in LISTED)) or any(short.startswith(t) for t in LISTED_COMM):
            import gzip
            fn = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_") + ".sass.gz"
            with gzip.open(os.path.join(OUT, "sass", fn), "wt") as f:
                f.write("Function : " + name + "\n" + body)
    ptx = subprocess.run(["cuobjdump", "-ptx", LIB], capture_output=True, text=True).stdout
    ptx_counts = {p: len(re.findall(re.escape(p), ptx)) for p in
                  ("tcgen05.mma", "tcgen05.ld", "tcgen05.alloc", "tcgen05.commit", "cp.async.bulk.tensor",
                   "multimem.ld_reduce", "multimem.st", "red.relaxed.sys.global.add.v4.f32", "ld.relaxed.sys.global.v4.f32",
                   "st.relaxed.sys.global.v4.f32", "atom.release.sys", "mbarrier.try_wait", "atom.relaxed.sys.global.add.v4.f32",
                   "fence.proxy.async", "griddepcontrol.wait", "st.release.sys.global.u32", "ld.acquire.sys.global.u32")}
    with open(os.path.join(OUT, "sass_summary.md"), "w") as f:
        f.write("# SASS evidence per kernel (cuobjdump -sass mpit_b200/lib/libmpit_kernels.so, sm_100a)\n\n")
        f.write("| kernel | instrs | " + " | ".join(k for k, _ in KEYS) + " |\n|---|---:|" + "---:|" * len(KEYS) + "\n")
        for short, n, cnt in sorted(rows):
            f.write(f"| `{short}` | {n} | " + " | ".join(str(cnt[k]) for k, _ in KEYS) + " |\n")
        f.write("\nPTX instruction counts (cuobjdump -ptx; empty if the library embeds no PTX):\n\n")
        for k, v in ptx_counts.items():
            f.write(f"- `{k}`: {v}\n")
    print(f"{len(rows)} kernels -> profiles/sass_summary.md")


if __name__ == "__main__":
    sys.exit(main())
