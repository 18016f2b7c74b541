#!/usr/bin/env python
"""Instruction histogram of the largest loop of one kernel in an object file / library (cuobjdump -sass).
    python tools/sass_loop.py <file> <mangled kernel name>"""
import re
import subprocess
import sys
from collections import Counter


def loop_hist(path, fun):
    txt = subprocess.run(["cuobjdump", "-sass", "-fun", fun, path], capture_output=True, text=True).stdout
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
    op = lambda t: re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0]
    whole = Counter(op(t) for _, t in ins)
    if best is None:
        return len(ins), whole, 0, Counter()
    n, tgt, a = best
    return len(ins), whole, n, Counter(op(t) for x, t in ins if tgt <= x <= a)


if __name__ == "__main__":
    total, whole, n, c = loop_hist(sys.argv[1], sys.argv[2])
    print("kernel: %d instructions (%d bytes); largest loop: %d instructions (%d bytes)" % (total, total * 16, n, n * 16))
    print("loop:", ", ".join("%s %d" % kv for kv in c.most_common(12)))
