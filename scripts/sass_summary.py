#!/usr/bin/env python
"""Per-kernel SASS evidence for profiles/sass_summary.txt (no GPU needed):

    python scripts/sass_summary.py > profiles/sass_summary.txt

For every kernel of madmom_b200/lib/libmadmom_b200.so: instruction count, registers / spills / shared memory
(cuobjdump -res-usage), and the counts of the mnemonics that show what the code is made of: UBLKCP (1-D bulk TMA
copy, cp.async.bulk), SYNCS (mbarrier), FADD2 / FMUL2 / FFMA2 (packed FP32 pairs), FFMA / FADD / FMUL, LDS / STS,
SHFL, MUFU, DADD / DMUL / DFMA (the float64 STFT), tensor-core ops (none expected: the path is FP32 FMA by design)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "madmom_b200", "lib", "libmadmom_b200.so")
WATCH = ["UBLKCP", "UTMALDG", "SYNCS", "FADD2", "FMUL2", "FFMA2", "FFMA", "FADD", "FMUL", "DFMA", "DADD", "DMUL", "LDS",
         "STS", "LDG", "STG", "SHFL", "MUFU", "BAR", "ATOMS", "HMMA", "UTCHMMA", "UTCMMA", "MULTIMEM", "ST.E.MULTIMEM"]


def demangle(names):
    out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out)) if len(out) == len(names) else {n: n for n in names}


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
    usage = {}
    cur = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        if cur and "REG:" in line:
            usage[cur] = " ".join(line.split())
            cur = None
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if cur and m:
            op = m.group(1)
            kernels[cur]["_total"] += 1
            kernels[cur][op.split(".")[0]] += 1
            if op.startswith("ST.E.MULTIMEM") or "MULTIMEM" in op:
                kernels[cur]["MULTIMEM"] += 1
    names = demangle(list(kernels))
    print("SASS summary of %s" % os.path.relpath(LIB, ROOT))
    print("architectures in the fat binary: %s" % ", ".join(arch))
    print("columns: kernel | instructions | resource usage | watched mnemonics (count)")
    print()
    for k, c in kernels.items():
        short = re.sub(r"\((?:int|bool)\)", "", names[k])      # cu++filt prints template values as casts
        short = re.sub(r"\([^()]*\)$", "", short)                  # drop the parameter list
        watched = " ".join("%s=%d" % (w, c[w]) for w in WATCH if c.get(w))
        print("%s\n    instr %d | %s\n    %s" % (short, c["_total"], usage.get(k, "?"), watched))
    tot = collections.Counter()
    for c in kernels.values():
        tot.update(c)
    print()
    print("whole library: " + " ".join("%s=%d" % (w, tot[w]) for w in WATCH if tot.get(w)))
    tensor = sum(tot[w] for w in ("HMMA", "UTCHMMA", "UTCMMA"))
    print("tensor-core instructions: %d (by design: BASELINE.json north_star asks for an FP32 FMA filterbank and the FFT "
          "is not a dense contraction)" % tensor)


if __name__ == "__main__":
    sys.exit(main())
