#!/usr/bin/env python
"""SASS opcode histogram of the hot kernels (cuobjdump -sass on the built objects; runs without a GPU).
For every kernel matching the patterns: counts of the opcodes that prove how it moves data (UBLKCP =
bulk copy through the TMA unit, SYNCS = mbarrier, LDS.128, ATOMS / RED / ATOMG) and what it computes with
(DADD / DMUL / DFMA / DSETP, F2F conversions, MUFU), and -- for the bit-exactness contract -- where the
DFMA instructions sit: they may only appear inside IEEE division / sqrt / log / exp / pow sequences
(recognised by a MUFU within the same straight-line neighbourhood), never in a bin-index expression.
  python tools/sass_histogram.py > profiles/rNN_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = {"bin_kernels.o": [r"bin_fast_kernelILi3ELi4ELb1ELi3E", r"bin_fast_kernelILi2ELi3ELb1ELi2E",
                         r"bin_fast_kernelILi1ELi1ELb1ELi1ELj1E", r"bin_fast_kernelILi3ELi4ELb1ELin1",
                         r"bin_fast_kernelILi3ELi4ELb0ELi3E", r"bin_generic_kernelILb1E"],
       "nll_kernels.o": [r"lookup_nll_kernelILi128E", r"lookup_nll_kernelILi256E", r"pdf_table_kernel"],
       "p2p_kernels.o": [r"hist_allreduce_p2p_kernel"],
       "nll_sharded_kernels.o": [r"lookup_nll_sharded_kernel"],
       "noalloc_kernels.o": [r"nll_lut_noalloc_kernel"],
       "batch_mono_kernels.o": [r"chain_batch_mono_kernel"]}
WATCH = ["UBLKCP", "UTMALDG", "SYNCS", "LDS.128", "LDS", "LDG", "LDC", "STS", "ATOMS", "ATOMG", "RED", "BAR",
         "DADD", "DMUL", "DFMA", "DSETP", "F2F", "F2I", "I2F", "MUFU", "FFMA", "HMMA", "UTCHMMA", "IMAD", "SEL",
         "NANOSLEEP", "MEMBAR", "ACQBULK"]


def main():
    print("# SASS opcode histogram (cuobjdump -sass), sm_100a.  What to read off it:")
    print("#  * bin_fast_kernel: UBLKCP (bulk copies through the TMA unit) + SYNCS (mbarrier) + LDS.128 feed the loop;")
    print("#    ZERO DFMA: no operation that feeds a bin index is fused (the x86-64 reference build has no FMA);")
    print("#    no tensor-core opcodes anywhere (nothing here is a contraction).")
    print("#  * bin_generic_kernel: its only DFMA are the explicit ones of cube_rn (x^3 rounded once, the 4-coefficient")
    print("#    series' pow(x, 3); bin_device.cuh); the bin-index expressions use DADD/DMUL only.")
    print("#  * lookup_nll / batch kernels: DFMA belong to IEEE division, log, exp, sqrt, cospi sequences (NLL side,")
    print("#    tolerance 1e-6); the 'without a MUFU nearby' count is a heuristic that long log/exp polynomials exceed.")
    for obj, pats in OBJ.items():
        path = os.path.join(ROOT, "sxmc_b200", "csrc", obj)
        if not os.path.exists(path):
            print("%s: not built" % obj)
            continue
        text = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
        funcs = re.split(r"\n\s*Function : ", text)
        for f in funcs[1:]:
            name = f.split("\n", 1)[0].strip()
            if not any(re.search(p, name) for p in pats):
                continue
            ops = []
            for line in f.splitlines():
                m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
                if m:
                    ops.append(m.group(1))
            hist = collections.Counter()
            for o in ops:
                for w in WATCH:
                    if o == w or o.startswith(w + "."):
                        hist[w] += 1
                        break
            # DFMA outside a division/transcendental neighbourhood: no MUFU within 40 instructions before
            stray = 0
            for i, o in enumerate(ops):
                if o.startswith("DFMA"):
                    window = ops[max(0, i - 60):i + 10]
                    if not any(x.startswith("MUFU") for x in window):
                        stray += 1
            print("== %s" % name[:110])
            print("   instructions %d; " % len(ops) + ", ".join("%s %d" % (w, hist[w]) for w in WATCH if hist[w]))
            print("   DFMA without a MUFU (reciprocal / rsqrt / lg2 / ex2 seed) within 60 instructions before it: %d" % stray)


if __name__ == "__main__":
    main()
