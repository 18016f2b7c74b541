"""gpurun_out/ (scratch) -> profiles/ (tracked): the launch list as captured, a per-kernel table of it, the DRAM traffic
per packet that bench.py's `roofline.traffic` reads, the bench lines, and the chosen metrics of the ncu --set full reports.
Run after tools/profile_round.sh:  python tools/summarize_profiles.py r1"""
import collections
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
           "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
           "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
           "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def short(name):
    return name.split("(")[0].replace("void ", "").strip()


def launch_table():
    src = os.path.join(G, "launches_full.csv")
    lines = [l for l in open(src) if l.startswith('"')]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    per = collections.OrderedDict()
    for r in rows:
        k = (r["ID"], short(r["Kernel Name"]), r["Grid Size"])
        per.setdefault(k, {})[r["Metric Name"]] = (float(r["Metric Value"].replace(",", "")), r["Metric Unit"])
    agg = collections.OrderedDict()
    for (_, name, grid), m in per.items():
        t, tu = m["gpu__time_duration.sum"]
        t_ms = t * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(tu, 1e-6)

        def nbytes(key):
            v, u = m[key]
            return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        a = agg.setdefault((name, grid), [0, 0.0, 0.0])
        a[0] += 1
        a[1] += t_ms
        a[2] += nbytes("dram__bytes_read.sum") + nbytes("dram__bytes_write.sum")
    shutil.copy(src, os.path.join(P, "%s_launches_full.csv" % tag))
    out = ["per-kernel summary of profiles/%s_launches_full.csv (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
           % tag, "--clock-control none -c 400, command: python bench.py --no-cpu --steps 2 --warmup 3; times are cold-cache and serialised)", "",
           "%-28s %-18s %8s %12s %14s" % ("kernel", "grid", "launches", "ms/launch", "DRAM GB/launch")]
    total = sum(a[1] for a in agg.values())
    for (name, grid), a in agg.items():
        out.append("%-28s %-18s %8d %12.3f %14.3f   %4.1f %% of the listed time" % (name, grid, a[0], a[1] / a[0], a[2] / a[0] / 1e9, 100 * a[1] / total))
    open(os.path.join(P, "%s_launch_summary.txt" % tag), "w").write("\n".join(out) + "\n")
    # traffic per packet for the three big kernels, from the resident-leg launches (grid of 113664 packets)
    traffic = {"source": "ncu dram__bytes_read.sum + dram__bytes_write.sum per launch, profiles/%s_launches_full.csv (113664 C4 packets per launch), "
                         "divided by the packets of the launch" % tag}
    npk = 113664
    alg = {"viterbi_kernel": 436800, "demod_kernel": 241920, "traceback_kernel": 266175.0}
    for (name, grid), a in agg.items():
        base = name.split("<")[0]
        if base in alg:
            g0 = int(grid.strip("()").split(",")[0])
            full = {"viterbi_kernel": npk // 32, "traceback_kernel": npk // 32, "demod_kernel": None}[base]
            if base == "demod_kernel":
                full = npk * 3 // 8  # 113664 packets x 3 work items (105 symbols in runs of 35) / 8 items per block
            if full is not None and g0 == full and base not in traffic:
                traffic[base] = {"dram_bytes_per_packet": a[2] / a[0] / npk, "algorithmic_bytes_per_packet": alg[base]}
    json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)
    return out


def ncu_report(rep, title):
    path = os.path.join(G, rep)
    if not os.path.exists(path):
        return ["(%s missing)" % rep]
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    out = [title, ""]
    for r in data:
        d = dict(zip(hdr, r))
        out.append("%s   grid %s block %s" % (short(d["Kernel Name"]), d.get("Grid Size", ""), d.get("Block Size", "")))
        for m in METRICS:
            if m in d:
                out.append("    %-70s %s %s" % (m, d[m], units[hdr.index(m)]))
        stalls = []
        for h in hdr:
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") or \
               h.startswith("smsp__average_warp_latency_issue_stalled_") and h.endswith(".ratio"):
                try:
                    stalls.append((float(d[h]), h.split("stalled_")[1].split("_per_issue")[0].replace(".ratio", "")))
                except ValueError:
                    pass
        stalls.sort(reverse=True)
        if stalls:
            out.append("    stalls (warps per issue-active cycle): " + ", ".join("%s %.2f" % (n, v) for v, n in stalls[:6]))
        out.append("")
    return out


if __name__ == "__main__":
    os.makedirs(P, exist_ok=True)
    launch_table()
    for f, dst in (("bench_default.json", "%s_bench_default.json" % tag), ("bench_reference.json", "%s_bench_reference.json" % tag)):
        if os.path.exists(os.path.join(G, f)):
            shutil.copy(os.path.join(G, f), os.path.join(P, dst))
    txt = ncu_report("prof_round.ncu-rep", "ncu --set full --clock-control none --import-source on -k regex:'viterbi|demod|traceback' -s 9 -c 3\n"
                     "command: python bench.py --no-cpu --steps 2 --warmup 3 --mpdu-len 400 --packets 113664   (short packets: same occupancy and\n"
                     "loops as the benchmark, 10x less memory for ncu to save and restore per pass; durations are for 11-symbol / 3270-step packets)")
    open(os.path.join(P, "%s_ncu_viterbi_demod_traceback.txt" % tag), "w").write("\n".join(txt) + "\n")
    txt = ncu_report("prof_front.ncu-rep", "ncu --set full --clock-control none --import-source on -k regex:'presiso|cfo_copy|lsig|sync_event|run_edges' -c 6\n"
                     "same command: the front-end kernels (SURVEY 8(f) N3 + N1), first launches of the front-end leg (28416 short packets, 45 M samples)")
    open(os.path.join(P, "%s_ncu_front_end.txt" % tag), "w").write("\n".join(txt) + "\n")
    txt = ncu_report("prof_demod2.ncu-rep", "ncu --set full --clock-control none --import-source on -k regex:demod -s 2 -c 1\n"
                     "command: python bench.py --resident-only --config C5 --packets 14208 --steps 2 --warmup 1   (the two-antenna demodulator on\n"
                     "full-size C5 packets: HT MCS15, 64 symbols, two streams)")
    for rep, kre in (("prof_round.ncu-rep", "demod"), ("prof_demod2.ncu-rep", "demod"), ("prof_round.ncu-rep", "viterbi")):
        if os.path.exists(os.path.join(G, rep)):
            txt += ["", "per opcode (tools/ncu_ops.py %s %s):" % (rep, kre)]
            txt += subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_ops.py"), os.path.join(G, rep), kre, "14"],
                                  capture_output=True, text=True).stdout.splitlines()
    open(os.path.join(P, "%s_ncu_demod2_and_opcodes.txt" % tag), "w").write("\n".join(txt) + "\n")
    print(open(os.path.join(P, "%s_launch_summary.txt" % tag)).read())
    print(open(os.path.join(P, "traffic.json")).read())
