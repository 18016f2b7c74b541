#!/usr/bin/env python
"""Evidence file for the Viterbi roofline argument: the SASS of the main loop of viterbi_kernel<rate> as built (first and
last lines of the loop body in full, the middle elided), its instruction histogram, and the counts per trellis step and
per add-compare-select.   python tools/sass_excerpt.py > profiles/r2_sass_viterbi_bodies.txt"""
import os
import re
import subprocess
import sys
from collections import Counter

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fork-gr-ieee80211_b200", "libw80211rx.so")
KERNELS = [("1/2", 0, 6), ("2/3", 1, 6), ("3/4", 2, 6), ("5/6", 3, 5)]


def loop_of(fun):
    txt = subprocess.run(["cuobjdump", "-sass", "-fun", fun, LIB], capture_output=True, text=True).stdout
    ins = []
    for l in txt.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    best = None
    for a, t in ins:
        m = re.search(r"BRA\s+(?:\S+,\s*)?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            tgt = int(m.group(1), 16)
            n = sum(1 for x, _ in ins if tgt <= x <= a)
            if best is None or n > best[0]:
                best = (n, tgt, a)
    _, tgt, a = best
    return len(ins), [(x, t) for x, t in ins if tgt <= x <= a]


def main():
    op = lambda t: re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0]
    print("SASS of the main loops of viterbi_kernel<CR> in fork-gr-ieee80211_b200/libw80211rx.so (cuobjdump -sass, sm_100a),")
    print("made by tools/sass_excerpt.py.  One loop body = `steps` trellis steps x 64 add-compare-selects for 32 packets (one per lane).")
    print()
    for rate, cr, steps in KERNELS:
        fun = "_Z14viterbi_kernelILi%dEEvPK4int4PKiPK11w80211_descPKfP5uint2i" % cr
        total, body = loop_of(fun)
        c = Counter(op(t) for _, t in body)
        n = len(body)
        acs = steps * 64
        print("== rate %s: kernel %d instructions, loop body %d instructions (%d bytes) for %d steps" % (rate, total, n, 16 * n, steps))
        print("   histogram: " + ", ".join("%s %d" % kv for kv in c.most_common()))
        print("   per add-compare-select: %.3f instructions (FADD %.3f, FMNMX %.3f, SHF %.3f, other %.3f)"
              % (n / acs, c["FADD"] / acs, c["FMNMX"] / acs, c["SHF"] / acs, (n - c["FADD"] - c["FMNMX"] - c["SHF"]) / acs))
        if rate == "3/4":
            print("   loop body, first 40 and last 24 instructions:")
            for x, t in body[:40]:
                print("     /*%04x*/  %s" % (x, t))
            print("     ... %d instructions of the same four kinds ..." % (n - 64))
            for x, t in body[-24:]:
                print("     /*%04x*/  %s" % (x, t))
        print()


if __name__ == "__main__":
    main()
