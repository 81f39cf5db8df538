#!/usr/bin/env python
"""Regenerate profiles/sass_mnemonics_per_kernel.txt: for every kernel of zoo/native/libzoo_cuda.so, the count of the SASS
instructions that prove which hardware path it uses (`cuobjdump -sass`; B200_PROFILING.md lists the mnemonics):
UTCHMMA / UTCQMMA = tcgen05.mma (bf16 / FP8 kind::f8f6f4), LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UTCATOMSWS = TMEM alloc,
UTMALDG / UTMASTG / UTMAREDG = TMA tensor load / store / reduce-add, SYNCS = mbarrier, LDGMC / STGMC / REDGMC... = multimem (NVLS),
HMMA = warp-level mma.sync, LDSM / MOVM = ldmatrix / movmatrix, FFMA2 = packed fp32 FMA."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "zoo", "native", "libzoo_cuda.so")
KEEP = re.compile(r"^(UTC|UTMA|LDTM|STTM|SYNCS|HMMA|QMMA|LDSM|MOVM|LDGMC|STGMC|REDGMC|ATOMGMC|MULTIMEM|FFMA2|UBLKCP|UBLKRED|CCTL\.IVALL|ATOMS\.CAST)")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
    counts, cur, k = collections.OrderedDict(), None, 0
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = re.sub(r"\(.*", "", names[k].replace("zoo::", "").replace("void ", ""))
            k += 1
            counts.setdefault(cur, collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
        if m and cur is not None and KEEP.match(m.group(1)):
            op = m.group(1)
            op = re.sub(r"\.(SP|2CTA)?\..*$", "", op) if op.startswith(("UTCHMMA", "UTCQMMA")) else op
            counts[cur][op] += 1
    out = ["# cuobjdump -sass zoo/native/libzoo_cuda.so : instructions per kernel that name the hardware path (count); regenerate with",
           "# python dev/sass_mnemonics.py.  UTCHMMA / UTCQMMA = tcgen05.mma bf16 / FP8, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit,",
           "# UTCATOMSWS = TMEM alloc / dealloc, UTMALDG / UTMASTG / UTMAREDG = TMA tensor load / store / reduce-add, SYNCS = mbarrier ops,",
           "# LDGMC / STGMC = multimem.ld_reduce / multimem.st (NVLS), HMMA / LDSM / MOVM = mma.sync / ldmatrix / movmatrix (the head kernel only)."]
    for kern, c in counts.items():
        for op in sorted(c):
            out.append(f"{kern} {op} {c[op]}")
    path = os.path.join(ROOT, "profiles", "sass_mnemonics_per_kernel.txt")
    with open(path, "w") as fh:
        fh.write("\n".join(out) + "\n")
    kernels = [k for k, c in counts.items() if c]
    print(f"{len(kernels)} kernels with listed instructions -> {path}")
    tc = sorted({k for k, c in counts.items() if any(o.startswith(("UTCHMMA", "UTCQMMA")) for o in c)})
    print("tcgen05.mma in:", ", ".join(tc))
    print("mma.sync in:", ", ".join(sorted(k for k, c in counts.items() if any(o.startswith("HMMA") for o in c))))
    print("multimem in:", ", ".join(sorted(k for k, c in counts.items() if any("GMC" in o for o in c))))


if __name__ == "__main__":
    sys.exit(main())
