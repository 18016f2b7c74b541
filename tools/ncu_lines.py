"""Attribute the warp-stall samples and executed instructions of one kernel in an `ncu --page source --csv` export to CUDA
source lines, through the line table of the object file (nvdisasm -g).
  python tools/ncu_lines.py <source.csv> <object.o> <kernel-name substring> [lines per bucket]"""
import collections, csv, os, re, subprocess, sys, tempfile
src, obj, kname = sys.argv[1:4]
bucket = int(sys.argv[4]) if len(sys.argv) > 4 else 1
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cub], cwd=tmp, capture_output=True, text=True).stdout
addr2line, cur, infn = {}, None, False
for l in dis.splitlines():
    if l.startswith(".text."):
        infn = kname in l
    if not infn:
        continue
    m = re.search(r'//## File ".*?/([^/"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1), int(m.group(2)) // bucket * bucket)
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+\S", l)
    if m:
        addr2line[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src)))
hdr = rows[1]
iS, iI, iA = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Address")
body = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name":  # a second launch of the kernel in the same export: the first one is enough
        break
    if len(r) > iS:
        body.append(r)
base = int(body[0][iA], 16)
tot = sum(float(r[iS] or 0) for r in body)
toti = sum(float(r[iI] or 0) for r in body)
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
agg = collections.defaultdict(lambda: [0.0, 0.0, collections.Counter()])
for r in body:
    k = addr2line.get(int(r[iA], 16) - base)
    a = agg[k]
    a[0] += float(r[iS] or 0)
    a[1] += float(r[iI] or 0)
    for i in stall:
        a[2][hdr[i][6:]] += float(r[i] or 0)
print("%d SASS lines, %.0f samples, %.0f warp-instructions" % (len(body), tot, toti))
for k, (s_, i_, st) in sorted(agg.items(), key=lambda x: -x[1][0])[:30]:
    print("%-34s samples %5.1f%%  inst %5.1f%%  %s" % (k, s_ / tot * 100, i_ / toti * 100, [(a, int(b)) for a, b in st.most_common(2)]))
