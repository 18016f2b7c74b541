"""Per-opcode view of one kernel in an ncu report: instructions, stall samples, shared-memory wavefronts (actual / ideal).
   python tools/ncu_ops.py <report.ncu-rep> <kernel regex> [top N]"""
import collections
import csv
import io
import subprocess
import sys

rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 24
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + kre], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
body = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break
    body.append(r)


def f(r, k):
    try:
        return float(r[ix[k]])
    except Exception:
        return 0.0


ts, ti = sum(f(r, "# Samples") for r in body), sum(f(r, "Instructions Executed") for r in body)
tw, twi = sum(f(r, "L1 Wavefronts Shared") for r in body), sum(f(r, "L1 Wavefronts Shared Ideal") for r in body)
print(rows[0][1][:100])
print("%d SASS lines, %d samples, %d warp-instructions, shared wavefronts %d (ideal %d)" % (len(body), ts, ti, tw, twi))
byop = collections.defaultdict(lambda: [0, 0, 0, 0])
for r in body:
    op = [x for x in r[ix["Source"]].split() if not x.startswith("@")][0]
    b = byop[op]
    b[0] += f(r, "Instructions Executed")
    b[1] += f(r, "# Samples")
    b[2] += f(r, "L1 Wavefronts Shared")
    b[3] += f(r, "L1 Wavefronts Shared Ideal")
for op, b in sorted(byop.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%-22s instr %10d (%4.1f%%)  samples %6d (%4.1f%%)  smem wavefronts %10d ideal %10d" % (op, b[0], 100 * b[0] / ti, b[1], 100 * b[1] / ts, b[2], b[3]))
