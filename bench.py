#!/usr/bin/env python
"""bench.py -- decoded PSDU throughput of the demod + decode hot path on B200.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload of `value` / `e2e` (BASELINE.json configs[3], the configuration the 50 Gbit/s target is quoted on):
C4 = VHT MCS8 (256-QAM r3/4), 1 spatial stream, long GI, APEP 4092 (one 4088-byte MPDU per A-MPDU), AWGN at 30 dB,
synthetic.  A step is one pass of the whole path (header, demod, Viterbi, traceback, descramble + FCS) over one batch
of packets.  `value` is measured with the batch already in HBM; `e2e` goes through the C ABI with pinned HOST buffers
holding fc32 samples (the item type of the reference's block API), copies inside the timed region.  The other
configurations of BASELINE.json (C2, C3, C5) are timed resident and reported under `configs`; `batch_curve` is the
resident throughput and single-call latency against batch size; `e2e_blocks` is the GNU Radio block path (gr/lib under
the mock scheduler).  One JSON line on stdout (rank 0).

Both arms decode the same packets: packets are generated in chunks of 32 whose seeds depend only on (config, rank,
chunk number), the GPU arm tiles chunks 0..15, the CPU arm's worker i decodes chunks i, i + cores, ...
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

TRACE = bool(os.environ.get("W80211_BENCH_TRACE"))  # per-call host and per-slot event times on stderr (diagnostics)
METRIC = "decoded_psdu_gbit_per_s"
UNIT = "Gbit/s"
CHUNK = 32          # packets per generated chunk (one seed each)
N_CHUNKS = 16       # chunks the GPU arm tiles: 512 distinct packets
N_UNIQUE = CHUNK * N_CHUNKS

# BASELINE.json configs 2..5 as SURVEY.md 8(d) makes them concrete.  payload = bytes counted per FCS-valid packet.
CONFIGS = {
    "C2": {"workload": "C2: legacy 54 Mb/s 64-QAM r3/4, 1500 B, AWGN 30 dB, 20 MHz", "fmt": "L", "mcs": 7, "nss": 1, "length": 1500,
           "short_gi": False, "payload": 1500, "idx": 2},
    "C3": {"workload": "C3: HT MCS7 64-QAM r5/6 1SS short-GI, 4095 B, AWGN 30 dB, 20 MHz", "fmt": "HT", "mcs": 7, "nss": 1, "length": 4095,
           "short_gi": True, "payload": 4095, "idx": 3},
    "C4": {"workload": "C4: VHT MCS8 256-QAM r3/4 1SS long-GI, APEP 4092 B (MPDU 4088 B), AWGN 30 dB, 20 MHz", "fmt": "VHT", "mcs": 8,
           "nss": 1, "length": 4088, "short_gi": False, "payload": 4088, "idx": 4},
    "C5": {"workload": "C5: HT MCS15 64-QAM r5/6 2SS 2x2 long-GI (demod2, zero forcing), 4095 B, AWGN 30 dB, 20 MHz", "fmt": "HT", "mcs": 15,
           "nss": 2, "length": 4095, "short_gi": False, "payload": 4095, "idx": 5},
    # tuning only (--resident-only): the two code rates no BASELINE configuration exercises
    "T12": {"workload": "T12 (tuning): VHT MCS3 16-QAM r1/2 1SS, MPDU 2040 B", "fmt": "VHT", "mcs": 3, "nss": 1, "length": 2040,
            "short_gi": False, "payload": 2040, "idx": 6},
    "T23": {"workload": "T23 (tuning): VHT MCS5 64-QAM r2/3 1SS, MPDU 4000 B", "fmt": "VHT", "mcs": 5, "nss": 1, "length": 4000,
            "short_gi": False, "payload": 4000, "idx": 7},
}
CONFIG = "C4"   # the configuration of `value`; --config changes it for --resident-only tuning runs
MPDU_LEN = 4088  # --mpdu-len: profiling variants of C4 only
WORKLOAD = CONFIGS["C4"]["workload"]


def gen_chunk(config, k, rank=0):
    """Chunk k of `config`'s packet sequence for `rank`: CHUNK packets, reproducible from (config, rank, k) alone."""
    import w80211rx
    gp, g = w80211rx.gen_ppdu, w80211rx.gen
    c = CONFIGS[config]
    fmt = {"L": g.F_L, "HT": g.F_HT, "VHT": g.F_VHT}[c["fmt"]]
    length = MPDU_LEN if config == "C4" else c["length"]
    seed = 0x80211 + 100000 * c["idx"] + 4096 * rank + k
    return gp.gen_batch(fmt, c["mcs"], c["nss"], length, CHUNK, snr_db=30, short_gi=c["short_gi"], seed=seed)


def payload_len(config):
    return MPDU_LEN if config == "C4" else CONFIGS[config]["payload"]


def msg_digest(msgs):
    """Order-independent digest of a set of published messages (bytes objects)."""
    h = 0
    for m in msgs:
        h ^= int.from_bytes(hashlib.sha256(m).digest()[:8], "little")
    return "%016x" % h


# ------------------------------------------------------------------ reference arm
def _ref_worker(args):
    """One oracle instance (the reference's blocks compiled unchanged) over chunks first, first + stride, ... of the
    packet sequence.  Only the two blocks' runs are timed (stream packing in Python is not the reference's work)."""
    config, first, stride, nchunks, rank = args
    from oracle import refgr
    import w80211rx
    gp = w80211rx.gen_ppdu
    dt, ok, n, digest_msgs = 0.0, 0, 0, []
    for k in range(first, first + stride * nchunks, stride):
        b = gen_chunk(config, k, rank)
        pk = gp.to_packets(b)
        streams = refgr.make_sig_stream(pk, nant=b["nrx"])
        t0 = time.perf_counter()
        llr = refgr.run_demod(streams)
        msgs = refgr.run_decode(llr)
        dt += time.perf_counter() - t0
        sent = {x.tobytes() for x in b["mpdu"]}  # a packet whose FCS fails publishes nothing: match by content, not by position
        ok += sum(1 for m in msgs if m[3:-1] in sent)
        n += CHUNK
        digest_msgs += msgs
    return dt, ok, n, msg_digest(digest_msgs), len(digest_msgs)


def _port_worker(args):
    """Fallback when oracle/_ref is missing: the plain-C restatement of the decode block (oracle/port) over LLRs made
    from C4 packets' coded bits (+1/-1 with Gaussian noise).  The demod half (< 1 % of the CPU path) is not in it."""
    config, first, stride, nchunks, rank = args
    import w80211rx
    from oracle import port
    g = w80211rx.gen
    npkt = CHUNK * nchunks
    rng = np.random.default_rng(first)
    mp = g.make_mpdus(npkt, MPDU_LEN, rng)
    psdu = g.vht_ampdu(mp)
    m = g.Mod(g.F_VHT, 8, 1, psdu.shape[1])
    _, bits = g.data_field_bits(m, psdu)
    coded = g.puncture(g.bcc_encode(bits), m.cr).astype(np.float32)
    llr = (2.0 * coded - 1.0) * 4.0 + rng.normal(0, 1.0, coded.shape).astype(np.float32)
    t0 = time.perf_counter()
    ok = 0
    for p in range(npkt):
        msgs, _ = port.decode_packet({"llr": llr[p], "format": 2, "mcs": 8, "len": psdu.shape[1], "cr": m.cr, "ampdu": 1,
                                      "trellis": m.trellis})
        ok += int(len(msgs) == 1 and msgs[0][3:-1] == mp[p].tobytes())
    return time.perf_counter() - t0, ok, npkt, None, 0


def reference_kind():
    """'reference': the reference's own blocks (oracle/_ref); 'port': the plain-C restatement of the decode block; None."""
    from oracle import port, refgr
    return "reference" if refgr.available() else ("port" if port.available() else None)


def run_reference_sample(config, cores, chunks_per_core, rank=0, first_chunk=0):
    """Worker i decodes chunks first_chunk + i, + cores, ... (chunks_per_core of them).
    -> dict(gbps over all cores, ok, n, slowest worker seconds, wall, digest of everything published, messages)."""
    import multiprocessing as mp
    kind = reference_kind()
    if kind == "reference":
        from oracle import refgr
        refgr.lib()  # map libref80211.so in this process too: the forked workers inherit it
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        t0 = time.perf_counter()
        worker = _ref_worker if kind == "reference" else _port_worker
        res = pool.map(worker, [(config, first_chunk + i, cores, chunks_per_core, rank) for i in range(cores)])
        wall = time.perf_counter() - t0
    ok = sum(r[1] for r in res)
    n = sum(r[2] for r in res)
    slowest = max(r[0] for r in res)  # block time only (generation and stream packing excluded), cores run side by side
    dig = None
    if kind == "reference":
        x = 0
        for r in res:
            x ^= int(r[3], 16)
        dig = "%016x" % x
    return {"gbps": ok * payload_len(config) * 8 / slowest / 1e9, "ok": ok, "n": n, "slowest": slowest, "wall": wall, "digest": dig,
            "nmsg": sum(r[4] for r in res)}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    kind = reference_kind()
    if kind is None:
        print(json.dumps({"impl": "reference", "unavailable": "neither oracle/_ref/libref80211.so nor oracle/libport80211.so is built (run __graft_entry__.build())"}))
        return 0
    cores = host_cores()
    cpc = max(1, args.ref_packets // CHUNK)  # chunks per core and step
    for _ in range(max(args.warmup, 0)):
        run_reference_sample(CONFIG, cores, 1)
    tot_ok, tot_n, tsum = 0, 0, 0.0
    for k in range(args.steps):
        # every step decodes the same chunks the GPU arm's batch is made of (0..15), then further ones of the same sequence
        r = run_reference_sample(CONFIG, cores, cpc)
        tot_ok += r["ok"]
        tot_n += r["n"]
        tsum += r["slowest"]
    value = tot_ok * payload_len(CONFIG) * 8 / tsum / 1e9
    what = "reference blocks (demod + decode) built unchanged, mock scheduler; only the blocks' run time is counted" if kind == "reference" else \
        "plain-C restatement of the decode block only (oracle/port; oracle/_ref missing), LLRs from the packets' coded bits"
    sample = "%d cores x %d packets per step, %d steps; %s" % (cores, cpc * CHUNK, args.steps, what)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tsum / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD},
            "sample_packets_per_step": cores * cpc * CHUNK,
            "packets_per_s": tot_ok / tsum, "fcs_ok": tot_ok, "packets": tot_n,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------ clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append([time.time()] + [x.strip() for x in ln.split(",")])

    def mark(self):
        return time.time()

    def stop(self, t_from=None, t_to=None):
        """Samples that arrived inside [t_from, t_to] (the timed region); all of them when no window is given."""
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                pass
        rows = [r[1:] for r in self.rows if len(r) >= 9 and (t_from is None or t_from <= r[0] <= t_to + 0.15)]
        sm = [float(r[0]) for r in rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        pw = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            for i, nm in enumerate(names):
                if r[4 + i].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w": float(np.median(pw)) if pw else None, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------ our arm: helpers
class Workload:
    """One configuration's batch: N_UNIQUE distinct packets (chunks 0..15 of the rank's sequence) tiled to P packets,
    resident in HBM; optionally the first PE packets also in pinned host memory (fc32, sc16, sc8)."""

    def __init__(self, torch, host, dev, config, P, rank, PE=0):
        self.config, self.P, self.PE = config, P, PE
        chunks = [gen_chunk(config, k, rank) for k in range(N_CHUNKS)]
        self.chunks = chunks
        iq_u, pk_u, ch_u = host.pack_batch(chunks)
        self.nant = len(iq_u)
        nsamp = self.nsamp = int(pk_u["nsamp"][0])
        reps = (P + N_UNIQUE - 1) // N_UNIQUE
        pk = np.zeros(P, host.PKT_IN_DTYPE)
        pk["iq_offset"] = np.arange(P, dtype=np.uint64) * nsamp
        pk["nsamp"] = nsamp
        pk["l_mcs"] = pk_u["l_mcs"][0]
        pk["l_len"] = pk_u["l_len"][0]
        chan = np.ascontiguousarray(np.tile(ch_u, (reps, 1))[:P])
        self.iq_dev = []
        for a in range(self.nant):
            t = torch.empty(P * nsamp, dtype=torch.complex64, device=dev)
            src = torch.from_numpy(iq_u[a]).to(dev)
            for r in range(reps):
                n = min(N_UNIQUE, P - r * N_UNIQUE) * nsamp
                t[r * N_UNIQUE * nsamp: r * N_UNIQUE * nsamp + n] = src[:n]
            self.iq_dev.append(t)
        self.pk_dev = torch.from_numpy(pk.view(np.uint8).copy()).to(dev)
        self.ch_dev = torch.from_numpy(chan.view(np.float32).reshape(-1).copy()).to(dev)
        self.bytes_cap = P * (payload_len(config) + 80) + 4096
        self.mp_cap = 2 * P + 64
        self.torch, self.host, self.dev = torch, host, dev
        if PE:
            self.iq_host = [t[:PE * nsamp].cpu().pin_memory() for t in self.iq_dev]
            self.pk_host = torch.from_numpy(pk[:PE].view(np.uint8).copy()).pin_memory()
            self.ch_host = torch.from_numpy(chan[:PE].view(np.float32).reshape(-1).copy()).pin_memory()
            reps_e = (PE + N_UNIQUE - 1) // N_UNIQUE
            self.q_host, self.pkq_host = {}, {}
            for bits in (16, 8):
                q_u, pkq_u, _ = host.quantize_sc16(iq_u, pk_u, bits=bits)
                self.q_host[bits] = [torch.from_numpy(np.tile(q, (reps_e, 1))[:PE * nsamp].copy()).pin_memory() for q in q_u]
                pkq = pk[:PE].copy()
                pkq["iq_scale"] = np.tile(pkq_u["iq_scale"], reps_e)[:PE]
                self.pkq_host[bits] = torch.from_numpy(pkq.view(np.uint8).copy()).pin_memory()

    def batch_in(self, npkt=None):
        """Device-resident input of the first npkt packets."""
        host = self.host
        n = self.P if npkt is None else npkt
        return host.Receiver.make_in([t.data_ptr() for t in self.iq_dev], self.pk_dev.data_ptr(), self.ch_dev.data_ptr(),
                                     mem=host.MEM_DEVICE, nsamples=n * self.nsamp, npkt=n)

    def batch_in_host(self, fmt):
        """Pinned host input of the first PE packets; fmt: 'fc32' | 'sc16' | 'sc8'."""
        host = self.host
        if fmt == "fc32":
            iq, pk, code = self.iq_host, self.pk_host, host.IQ_FC32
        else:
            bits = 16 if fmt == "sc16" else 8
            iq, pk, code = self.q_host[bits], self.pkq_host[bits], (host.IQ_SC16 if bits == 16 else host.IQ_SC8)
        bi = host.Receiver.make_in([t.data_ptr() for t in iq], pk.data_ptr(), self.ch_host.data_ptr(), mem=host.MEM_DEVICE,
                                   nsamples=self.PE * self.nsamp, npkt=self.PE, iq_format=code)
        bi.mem = host.MEM_HOST
        return bi

    def h2d_bytes(self, fmt):
        per = {"fc32": 8, "sc16": 4, "sc8": 2}[fmt]
        return self.PE * self.nsamp * per * self.nant + self.pk_host.numel() + self.ch_host.numel() * 4

    def dev_out(self):
        torch, host = self.torch, self.host
        t = {"desc": torch.empty(self.P * host.DESC_DTYPE.itemsize, dtype=torch.uint8, device=self.dev),
             "bytes": torch.empty(self.bytes_cap, dtype=torch.uint8, device=self.dev),
             "psdu_off": torch.empty(self.P, dtype=torch.int32, device=self.dev),
             "mpdus": torch.empty(self.mp_cap * 16, dtype=torch.uint8, device=self.dev)}
        bo = host.BatchOut()
        bo.mem = host.MEM_DEVICE
        bo.desc, bo.bytes, bo.bytes_capacity = t["desc"].data_ptr(), t["bytes"].data_ptr(), self.bytes_cap
        bo.psdu_off, bo.mpdus, bo.mpdu_capacity = t["psdu_off"].data_ptr(), t["mpdus"].data_ptr(), self.mp_cap
        return bo, t

    def host_out(self):
        torch, host = self.torch, self.host
        PE = self.PE
        bcap, mcap = PE * (payload_len(self.config) + 80) + 4096, 2 * PE + 64
        t = {"desc": torch.empty(PE * host.DESC_DTYPE.itemsize, dtype=torch.uint8).pin_memory(),
             "bytes": torch.empty(bcap, dtype=torch.uint8).pin_memory(),
             "psdu_off": torch.empty(PE, dtype=torch.int32).pin_memory(),
             "mpdus": torch.empty(mcap * 16, dtype=torch.uint8).pin_memory()}
        bo = host.BatchOut()
        bo.mem = host.MEM_HOST
        bo.desc, bo.bytes, bo.bytes_capacity = t["desc"].data_ptr(), t["bytes"].data_ptr(), bcap
        bo.psdu_off, bo.mpdus, bo.mpdu_capacity = t["psdu_off"].data_ptr(), t["mpdus"].data_ptr(), mcap
        return bo, t


STAGE_KEYS = ("header", "demod", "viterbi", "traceback", "post", "total", "host_plan", "host_collect")


def run_steps(rx, nsteps, bi, bos, depth=2, stage=None, counters=None):
    """nsteps batches through w80211_rx_submit / collect with `depth` slots in flight: the host plans batch i+1 (it needs
    the header kernel's descriptors) while the GPU decodes batch i.  -> (FCS-valid PSDU bits, FCS-valid MPDUs)"""
    L, h = rx.L, rx.h
    tot_bits = tot_ok = 0
    pend = []

    def collect(ps):
        nonlocal tot_bits, tot_ok
        tc0 = time.perf_counter()
        rx._check(L.w80211_rx_collect(h, ps))
        if TRACE:
            t = rx.timing(ps)
            print("trace collect slot %d wait %.2f ms | h2d %.2f hdr %.2f demod %.2f vit %.2f tb %.2f post %.2f d2h %.2f total %.2f plan %.2f coll %.2f"
                  % (ps, (time.perf_counter() - tc0) * 1e3, t.h2d, t.header, t.demod, t.viterbi, t.traceback, t.post, t.d2h, t.total,
                     t.host_plan, t.host_collect), file=sys.stderr)
        tot_bits += bos[ps][0].psdu_bits_ok
        tot_ok += bos[ps][0].n_crc_ok
        if stage is not None:
            t = rx.timing(ps)
            if counters is not None:
                counters["launches"] += t.launches
            for k in stage:
                stage[k] += getattr(t, k)

    for i in range(nsteps):
        s = i % depth
        if len(pend) == depth:
            collect(pend.pop(0))
        ts0 = time.perf_counter()
        rx._check(L.w80211_rx_submit(h, s, C.byref(bi), C.byref(bos[s][0])))
        if TRACE:
            print("trace submit slot %d took %.2f ms" % (s, (time.perf_counter() - ts0) * 1e3), file=sys.stderr)
        pend.append(s)
    for ps in pend:
        collect(ps)
    return tot_bits, tot_ok


def timed(torch, barrier, fn):
    """fn() between two CUDA events on the current stream, barrier + synchronize on both sides -> (result, ms)."""
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r = fn()
    barrier()
    e1.record()
    e1.synchronize()
    return r, e0.elapsed_time(e1)


def rooflines(m, nant, P, stage_ms, nsm, sm_mhz, hbm_peak):
    """Per-kernel rooflines of one configuration from its per-kernel CUDA-event times (ms per launch over P packets).
    m: gen.Mod of the packets.  Algorithmic work per unit as SURVEY.md 8(d) defines it:
      Viterbi   64 ACS per trellis step, 1 ACS = 3 lane-ops; peak = SMs x 128 lanes x SM clock / 3
      demod     per symbol nRx * nSymSamp * 8 B of IQ read + nCBPS * 4 B of LLRs written
      traceback 8 B of survivors read + 1 bit written per trellis step"""
    steps = float(m.trellis) * P
    out = {}
    if stage_ms.get("viterbi", 0) > 0:
        acs = steps * 64 / (stage_ms["viterbi"] * 1e-3)
        peak = nsm * 128 * sm_mhz * 1e6 / 3.0
        out["viterbi"] = {"bound": "fp32-issue", "achieved": acs / 1e12, "peak": peak / 1e12, "unit": "T ACS/s", "frac": acs / peak,
                          "ms": stage_ms["viterbi"]}
    if stage_ms.get("demod", 0) > 0:
        b = float(P) * m.nsym * (nant * m.nsym_samp * 8 + m.ncbps * 4)
        gbs = b / (stage_ms["demod"] * 1e-3) / 1e9
        out["demod"] = {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "ms": stage_ms["demod"],
                        "algorithmic_bytes": b}
    if stage_ms.get("traceback", 0) > 0:
        b = steps * 8.0 + steps / 8.0
        gbs = b / (stage_ms["traceback"] * 1e-3) / 1e9
        out["traceback"] = {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "ms": stage_ms["traceback"],
                            "algorithmic_bytes": b}
    return out


# ------------------------------------------------------------------ our arm
def main_gpu(args):
    import torch
    import w80211rx
    host = w80211rx.host
    if args.lib:
        if not args.resident_only:
            raise SystemExit("bench.py: --lib (a kernel-variant build) is for --resident-only tuning runs; the benchmark loads the product library")
        host.LIB_PATH = args.lib
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    lws = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    ndev = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if 1 < lws < ndev and ndev % lws == 0 and os.environ.get("W80211_BENCH_PACK_GPUS") != "1":
        # fewer ranks than GPUs on the box: spread them over the GPUs (0, 2, 4, 6 of eight) rather than packing them on the first
        # ones -- the host-fed leg is bound by the PCIe domains the GPUs hang off (profiles/r1_host_feed_8gpu.txt: GPUs 0-3 share
        # 93 GB/s, GPUs 4-7 141 GB/s), and the resident leg does not care
        local = local * (ndev // lws)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    nsm = torch.cuda.get_device_properties(local).multi_processor_count
    # two full waves of the thread-per-packet Viterbi kernel: 12 one-warp blocks per SM, one packet per lane
    P = args.packets if args.packets > 0 else nsm * 12 * 32 * 2
    PE = min(P, args.e2e_packets if args.e2e_packets > 0 else nsm * 12 * 32)  # host-fed legs: one full Viterbi wave per batch
    E2E_DEPTH = args.e2e_depth
    K = args.steps
    wl = Workload(torch, host, dev, CONFIG, P, rank, PE=0 if args.resident_only else PE)
    mod = wl.chunks[0]["mod"]
    rx = host.Receiver(local, nslots=max(2, E2E_DEPTH))
    bo_d = [wl.dev_out() for _ in range(2)]
    bi_dev = wl.batch_in()

    # ---- resident leg: inputs already in HBM, results left in HBM, two slots in flight
    clocks = ClockSampler(local)
    clocks.start()
    run_steps(rx, max(args.warmup, 3), bi_dev, bo_d)
    barrier()
    t_from = clocks.mark()
    (bits, npk_ok), ms = timed(torch, barrier, lambda: run_steps(rx, K, bi_dev, bo_d))
    # per-kernel durations (CUDA events on the slot's own stream): the same steps once more with ONE slot in
    # flight, because with two streams the events of one slot also span the other slot's kernels
    KSEQ = min(K, 5)
    stage = {k: 0.0 for k in STAGE_KEYS}
    counters = {"launches": 0}
    run_steps(rx, KSEQ, bi_dev, bo_d, depth=1, stage=stage, counters=counters)
    torch.cuda.synchronize()
    t_to = clocks.mark()
    launches = counters["launches"] * K // KSEQ
    stage_ms = {k: v / KSEQ for k, v in stage.items()}
    if args.resident_only:
        print(json.dumps({"config": CONFIG, "value": bits / (ms * 1e-3) / 1e9, "ms_per_step": ms / K, "packets": P, "lib": args.lib,
                          "fcs_ok_per_step": npk_ok / K, "clocks": clocks.stop(t_from, t_to),
                          "stage_ms_per_step": {k: round(v, 4) for k, v in stage_ms.items()}}))
        rx.close()
        if dist is not None:
            dist.destroy_process_group()
        return 0

    # ---- end-to-end legs: pinned host buffers -> C ABI -> pinned host results, E2E_DEPTH slots in flight
    bo_h = [wl.host_out() for _ in range(max(2, E2E_DEPTH))]

    def e2e_leg(fmt):
        bi = wl.batch_in_host(fmt)
        run_steps(rx, E2E_DEPTH, bi, bo_h, depth=E2E_DEPTH)
        (nbits, _), t = timed(torch, barrier, lambda: run_steps(rx, K, bi, bo_h, depth=E2E_DEPTH))
        return nbits, t

    e2e32_bits, e2e32_ms = e2e_leg("fc32")   # fc32 samples on the wire (8 B per sample): the reference API's item type, the headline
    e2e8_bits, e2e8_ms = e2e_leg("sc8")      # sc8 samples on the wire (2 B per sample)
    e2e16_bits, e2e16_ms = e2e_leg("sc16")   # sc16 samples on the wire (4 B per sample)
    clk = clocks.stop(t_from, t_to)
    d2h = int(bo_h[0][0].n_bytes) + int(bo_h[0][0].n_mpdu) * 16 + 64 + PE * host.DESC_DTYPE.itemsize

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    try:
        traffic_tab = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        traffic_tab = {}
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured copy bandwidth (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s (B200_PROFILING.md)"
    sm_clk = clk["sm_mhz"] or float(peaks.get("sm_max_mhz", 1965.0))

    extras = {}
    if world == 1:
        extras["batch_curve"] = batch_curve_leg(torch, rx, wl, bo_d, K)
        extras["signal_stage"] = signal_leg(torch, host, rx, wl, bo_d, dev, nsm, K)
        del bo_d, bi_dev
        wl.iq_dev = None
        torch.cuda.empty_cache()
        extras["configs"] = {}
        for cfg in ("C2", "C3", "C5"):
            extras["configs"][cfg] = config_leg(torch, host, rx, dev, cfg, P, rank, K, nsm, sm_clk, hbm_peak)
        extras["e2e_blocks"] = blocks_leg(args)

    # ---- max over ranks, totals over ranks
    vals = torch.tensor([ms, e2e16_ms, e2e32_ms, e2e8_ms], dtype=torch.float64, device=dev)
    sums = torch.tensor([float(bits), float(npk_ok), float(e2e16_bits), float(launches), float(e2e32_bits), float(e2e8_bits)],
                        dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    ms, e2e16_ms, e2e32_ms, e2e8_ms = vals.tolist()
    bits, npk_ok, e2e16_bits, launches, e2e32_bits, e2e8_bits = sums.tolist()
    if rank != 0:
        rx.close()
        if dist is not None:
            dist.destroy_process_group()
        return 0

    roofs = rooflines(mod, wl.nant, P, stage_ms, nsm, sm_clk, hbm_peak)
    rv = roofs.get("viterbi", {})
    # HBM view of the same kernel (secondary): LLRs read (total * 4 / trellis bytes per step) + one 64-bit survivor word written
    total = mod.nsym * mod.ncbps
    vit_bytes = float(mod.trellis) * P * (total * 4.0 / mod.trellis + 8.0)
    vit_gbs = vit_bytes / (stage_ms["viterbi"] * 1e-3) / 1e9 if stage_ms["viterbi"] > 0 else 0.0
    vit_traffic = (traffic_tab["viterbi_kernel"]["dram_bytes_per_packet"] * P) if MPDU_LEN == 4088 and CONFIG == "C4" and "viterbi_kernel" in traffic_tab else None

    def e2e_obj(nbits, t, fmt, note):
        return {"value": nbits / (t * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": wl.h2d_bytes(fmt), "d2h_bytes_per_step": d2h,
                "ms_per_step": t / K, "packets_per_step_per_gpu": PE, "slots_in_flight": E2E_DEPTH,
                "h2d_gb_per_s_per_gpu": wl.h2d_bytes(fmt) / (t / K * 1e-3) / 1e9, "iq_wire_format": note}

    line = {
        "metric": METRIC, "value": bits / (ms * 1e-3) / 1e9, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": CONFIGS[CONFIG]["workload"] if MPDU_LEN == 4088 else "PROFILING VARIANT, MPDU %d B, not the benchmark workload" % MPDU_LEN},
        "run": {"packets_per_step_per_gpu": P, "unique_packets": N_UNIQUE, "slots_in_flight": 2,
                "l2": "inputs larger than L2 (%.1f GB of IQ per step)" % (P * wl.nsamp * 8 * wl.nant / 1e9),
                "parallelism": "packets sharded by rank, no collective",
                "gpus": "rank r on GPU r * %d (%d ranks spread over the box's %d GPUs)" % (ndev // lws, lws, ndev) if 1 < lws < ndev and ndev % lws == 0 and os.environ.get("W80211_BENCH_PACK_GPUS") != "1" else "rank r on GPU r"},
        "packets_per_s": npk_ok / (ms * 1e-3), "fcs_ok_per_step": npk_ok / K / world,
        "e2e": e2e_obj(e2e32_bits, e2e32_ms, "fc32", "fc32 (gr_complex), the item type of the reference's block API; pinned host buffers through "
                       "w80211_rx_submit/collect, PCIe-bound: 70.4 KB of samples per 4088-byte MPDU"),
        "e2e_sc16": e2e_obj(e2e16_bits, e2e16_ms, "sc16", "sc16 + one scale per packet (SURVEY.md 8(f) N2, the USRP's native wire format), expanded to fc32 "
                            "on the GPU; the samples were quantised before the timed region: a figure for a caller that receives sc16, not for one holding fc32"),
        "e2e_sc8": e2e_obj(e2e8_bits, e2e8_ms, "sc8", "sc8 + one scale per packet (UHD otw_format=sc8), quantised before the timed region; only FCS-valid bits are counted"),
        "gpu_launches": int(launches),
        "clocks": clk,
        "stage_ms_per_step": stage_ms,
        "stage_note": "per-kernel CUDA-event durations from %d steps with one slot in flight, run right after the timed region" % KSEQ,
        "roofline": {"bound": "fp32-issue", "kernel": "viterbi_kernel (one thread per packet)", "achieved": rv.get("achieved"), "peak": rv.get("peak"),
                     "unit": "T ACS/s", "frac": rv.get("frac"), "traffic": vit_traffic, "traffic_source": traffic_tab.get("source"),
                     "peak_def": "SMs x 128 FP32 lanes x SM clock under load / 3 lane-ops per add-compare-select (SURVEY.md 8(d)); 64 ACS per trellis step",
                     "sm_mhz": sm_clk, "share_of_step": stage_ms["viterbi"] / (ms / K) if ms > 0 else None},
        "roofline_hbm": {"kernel": "viterbi_kernel", "bound": "hbm", "achieved": vit_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": vit_gbs / hbm_peak,
                         "algorithmic_bytes": vit_bytes, "peak_source": peak_src, "note": "secondary view: the kernel is issue bound, not HBM bound"},
        "roofline_demod": dict(roofs.get("demod", {}), kernel="demod_kernel", peak_source=peak_src),
        "roofline_traceback": dict(roofs.get("traceback", {}), kernel="traceback_kernel", peak_source=peak_src),
    }
    line.update(extras)
    # ---- CPU baseline beside it (rank 0, N = 1 only): the reference's own blocks on the host cores, over the first chunks of
    # the same packet sequence; the GPU decodes exactly those chunks once more through host buffers and the two sets of
    # published messages are compared
    if world == 1 and not args.no_cpu:
        kind = reference_kind()
        if kind is not None:
            cores = host_cores()
            cpc = max(1, args.cpu_packets // CHUNK)
            r = run_reference_sample(CONFIG, cores, cpc, rank)
            same = None
            if kind == "reference":
                cmp_chunks = [wl.chunks[k] if k < N_CHUNKS else gen_chunk(CONFIG, k, rank) for k in range(cores * cpc)]
                iq, pkts, chan = host.pack_batch(cmp_chunks)
                res = rx.rx(iq, pkts, chan)
                mine = res.messages()
                same = {"packets": len(pkts), "gpu_published": len(mine), "cpu_published": r["nmsg"], "sets_equal": msg_digest(mine) == r["digest"]}
            line["cpu_baseline"] = {"value": r["gbps"], "unit": UNIT, "cores": cores, "kind": kind,
                                    "sample": "chunks 0..%d (%d packets per core) of the same packet sequence on %d cores, %.1f s; %s; %d/%d FCS ok"
                                              % (cores * cpc - 1, cpc * CHUNK, cores, r["slowest"],
                                                 "reference demod + decode blocks built unchanged (oracle/_ref), mock scheduler, block run time only"
                                                 if kind == "reference" else "plain-C restatement of the decode block (oracle/port), LLRs from coded bits", r["ok"], r["n"]),
                                    "same_packets": same}
    print(json.dumps(line))
    rx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def batch_curve_leg(torch, rx, wl, bo_d, K):
    """Resident throughput (two slots in flight) and the latency of one call against batch size, Viterbi kernels picked by
    the library (one warp per packet below W80211_WARP_MAX_PKTS packets of a code rate, one thread per packet above)."""
    L, h = rx.L, rx.h
    out = []
    for n in (1, 32, 2048, 14208, 56832, 113664):
        if n > wl.P:
            continue
        bi = wl.batch_in(n)
        reps = max(4, min(64, (4 * 2048) // n)) if n < 14208 else max(4, min(K, 6))
        run_steps(rx, 3, bi, bo_d)
        (nbits, nok), t = timed(torch, torch.cuda.synchronize, lambda: run_steps(rx, reps, bi, bo_d))
        lat, st = [], {k: 0.0 for k in ("header", "demod", "viterbi", "traceback", "post", "total", "host_plan")}
        for _ in range(5):
            t0 = time.perf_counter()
            rx._check(L.w80211_rx_submit(h, 0, C.byref(bi), C.byref(bo_d[0][0])))
            rx._check(L.w80211_rx_collect(h, 0))
            lat.append((time.perf_counter() - t0) * 1e3)
            tm = rx.timing(0)
            for k in st:
                st[k] += getattr(tm, k) / 5
        out.append({"packets": n, "gbit_s": nbits / (t * 1e-3) / 1e9, "ms_per_batch": t / reps, "fcs_ok_per_batch": nok / reps,
                    "latency_ms_one_call": float(np.median(lat)), "stage_ms": {k: round(v, 4) for k, v in st.items()}})
    return out


def config_leg(torch, host, rx, dev, cfg, P, rank, K, nsm, sm_clk, hbm_peak):
    """Another configuration of BASELINE.json, resident: Gbit/s with two slots in flight, per-kernel times with one."""
    wl = Workload(torch, host, dev, cfg, P, rank)
    bos = [wl.dev_out() for _ in range(2)]
    bi = wl.batch_in()
    k2 = max(2, min(K, 6))
    run_steps(rx, 3, bi, bos)
    (nbits, nok), t = timed(torch, torch.cuda.synchronize, lambda: run_steps(rx, k2, bi, bos))
    if t < 200.0 and K > k2:
        # short steps (C2: 17 ms): the first step's header stage and host plan, which nothing hides, weigh too much in six
        k2 = K
        (nbits, nok), t = timed(torch, torch.cuda.synchronize, lambda: run_steps(rx, k2, bi, bos))
    stage = {k: 0.0 for k in STAGE_KEYS}
    run_steps(rx, 3, bi, bos, depth=1, stage=stage)
    st = {k: v / 3 for k, v in stage.items()}
    m = wl.chunks[0]["mod"]
    r = {"workload": CONFIGS[cfg]["workload"], "value": nbits / (t * 1e-3) / 1e9, "unit": UNIT, "packets_per_step": P, "ms_per_step": t / k2,
         "fcs_ok_per_step": nok / k2, "packets_per_s": nok / (t * 1e-3), "stage_ms_per_step": {k: round(v, 4) for k, v in st.items()},
         "viterbi_ms": st["viterbi"], "trellis": int(m.trellis), "code_rate": ["1/2", "2/3", "3/4", "5/6"][m.cr], "antennas": wl.nant,
         "rooflines": rooflines(m, wl.nant, P, st, nsm, sm_clk, hbm_peak)}
    del bos, bi, wl
    torch.cuda.empty_cache()
    return r


def signal_leg(torch, host, rx, wl, bo_d, dev, nsm, K):
    """Front end (SURVEY.md 8(f) N3 + N1) on resident raw samples: every packet's whole PPDU with its own carrier offset,
    320 empty samples between packets.  presiso -> detect (trigger + sync) -> signal (L-SIG check, CFO-corrected copy), all in
    HBM; the packets it hands over are then demodulated and decoded once."""
    L, h = rx.L, rx.h
    ppdu = np.concatenate([c["ppdu"][0] for c in wl.chunks])
    sync_idx = int(wl.chunks[0]["sync_idx"])
    nsamp = wl.nsamp
    P = min(wl.P, nsm * 12 * 32 // 2)  # a quarter of the resident batch: these stages scale linearly
    reps = (P + N_UNIQUE - 1) // N_UNIQUE
    npp, gap = ppdu.shape[1], 320
    stride = npp + gap
    rads = ((np.arange(N_UNIQUE) % 7 - 3) * 0.003).astype(np.float32)
    u = np.zeros((N_UNIQUE, stride), np.complex64)
    u[:, gap:] = (ppdu.astype(np.complex128) * np.exp(-1j * rads[:, None].astype(np.float64) * np.arange(npp)[None, :])).astype(np.complex64)
    nraw = P * stride
    raw_dev = torch.empty(nraw, dtype=torch.complex64, device=dev)
    usrc = torch.from_numpy(u.reshape(-1)).to(dev)
    for r in range(reps):
        n = min(N_UNIQUE, P - r * N_UNIQUE) * stride
        raw_dev[r * N_UNIQUE * stride: r * N_UNIQUE * stride + n] = usrc[:n]
    ac_dev = torch.empty(nraw, dtype=torch.float32, device=dev)
    cj_dev = torch.empty(nraw, dtype=torch.complex64, device=dev)
    KS = min(K, 5)
    torch.cuda.synchronize()
    for _ in range(2):
        rx._check(L.w80211_presiso_batch(h, raw_dev.data_ptr(), nraw, host.MEM_DEVICE, ac_dev.data_ptr(), cj_dev.data_ptr()))
    t0 = time.perf_counter()
    for _ in range(KS):
        rx._check(L.w80211_presiso_batch(h, raw_dev.data_ptr(), nraw, host.MEM_DEVICE, ac_dev.data_ptr(), cj_dev.data_ptr()))
    pre_ms = (time.perf_counter() - t0) * 1e3 / KS
    cap = 3 * P + 64
    sy = np.zeros(cap, host.SYNC_DTYPE)
    di, do = host.DetectIn(), host.DetectOut()
    di.mem, di.final, di.nsamples = host.MEM_DEVICE, 1, nraw
    di.preac, di.preconj, di.sig = ac_dev.data_ptr(), cj_dev.data_ptr(), raw_dev.data_ptr()
    do.sync_capacity, do.syncs = cap, sy.ctypes.data
    for _ in range(2):
        rx._check(L.w80211_detect_batch(h, C.byref(di), C.byref(do)))
    t0 = time.perf_counter()
    for _ in range(KS):
        rx._check(L.w80211_detect_batch(h, C.byref(di), C.byref(do)))
    det_ms = (time.perf_counter() - t0) * 1e3 / KS
    nflag, ntrig = int(do.nsync), int(do.n_trigger)
    true_pos = np.arange(P, dtype=np.int64) * stride + gap + sync_idx
    found = sy[:nflag]
    fpos = found["pos"].astype(np.int64)
    near = np.abs(fpos[np.searchsorted(fpos, true_pos - 4).clip(0, max(nflag - 1, 0))] - true_pos) <= 4 if nflag else np.zeros(P, bool)
    out_iq = torch.empty(P * nsamp, dtype=torch.complex64, device=dev)
    out_ch = torch.empty(max(nflag, 1) * 64, dtype=torch.complex64, device=dev)
    pk_out = np.zeros(max(nflag, 1), host.PKT_IN_DTYPE)
    si, so = host.SignalIn(), host.SignalOut()
    si.nsync, si.nant, si.mem, si.final = nflag, 1, host.MEM_DEVICE, 1
    si.iq[0], si.nsamples, si.syncs, si.start = raw_dev.data_ptr(), nraw, found.ctypes.data, 0
    so.mem, so.iq_capacity, so.pkts, so.chan = host.MEM_DEVICE, P * nsamp, pk_out.ctypes.data, out_ch.data_ptr()
    so.iq[0] = out_iq.data_ptr()
    for _ in range(2):
        rx._check(L.w80211_signal_batch(h, C.byref(si), C.byref(so)))
    lsig = copy = 0.0
    t0 = time.perf_counter()
    for _ in range(KS):
        rx._check(L.w80211_signal_batch(h, C.byref(si), C.byref(so)))
        lsig += so.lsig_ms
        copy += so.copy_ms
    wall = (time.perf_counter() - t0) * 1e3 / KS
    npk_sig = int(so.npkt)
    pk2 = torch.from_numpy(pk_out[:npk_sig].view(np.uint8).copy()).to(dev)
    bi2 = host.Receiver.make_in([out_iq.data_ptr()], pk2.data_ptr(), out_ch.data_ptr(), mem=host.MEM_DEVICE,
                                nsamples=int(so.iq_used), npkt=npk_sig)
    _, ok2 = run_steps(rx, 1, bi2, bo_d)
    copy_bytes = float(npk_sig) * (nsamp * 16.0 + 1024.0)
    return {"packets_per_step": P, "raw_samples_per_step": nraw,
            "presiso_ms_host_wall": pre_ms, "presiso_gbs": nraw * 20.0 / (pre_ms * 1e-3) / 1e9,
            "detect_ms_host_wall": det_ms, "triggers": ntrig, "sync_flags": nflag, "true_packets_flagged": int(near.sum()),
            "signal_ms_host_wall": wall, "lsig_kernel_ms": lsig / KS, "cfo_copy_kernel_ms": copy / KS,
            "cfo_copy_gbs": copy_bytes / (copy / KS * 1e-3) / 1e9 if copy > 0 else None,
            "accepted": npk_sig, "fcs_ok_after_demod_decode": int(ok2),
            "note": "w80211_presiso_batch + w80211_detect_batch + w80211_signal_batch, raw samples and results in HBM; not part of "
                    "`value` (the north-star path starts at demod's input)"}


def blocks_leg(args):
    """The GNU Radio block path (gr/lib demod -> decode streamed by the mock scheduler's chain runner over fc32 items in host
    memory) beside the reference's blocks under the same runner: tools/blocks_bench.py."""
    try:
        import tools.blocks_bench as bb
    except Exception:
        return None
    try:
        return bb.run(npkt=args.blocks_packets)
    except Exception as e:  # the harness is test infrastructure: its absence must not take the benchmark down
        return {"unavailable": "%s: %s" % (type(e).__name__, e)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--packets", type=int, default=int(os.environ.get("W80211_BENCH_PACKETS", "0")),
                    help="packets per step and GPU (0 = two full waves of the Viterbi kernel: SMs x 12 x 32 x 2)")
    ap.add_argument("--e2e-packets", type=int, default=0, help="packets per step and GPU of the host-fed legs (0 = SMs x 12 x 32)")
    ap.add_argument("--e2e-depth", type=int, default=3, help="slots in flight of the host-fed leg (2..4)")
    ap.add_argument("--mpdu-len", type=int, default=4088, help="profiling only: shorter C4 packets; the reported workload is the default")
    ap.add_argument("--ref-packets", type=int, default=512, help="--impl reference: packets per core and step")
    ap.add_argument("--cpu-packets", type=int, default=128, help="packets per core of the cpu_baseline sample inside the GPU arm")
    ap.add_argument("--blocks-packets", type=int, default=8192, help="packets of the e2e_blocks leg")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--config", default="C4", choices=["C2", "C3", "C4", "C5", "T12", "T23"],
                    help="tuning runs (with --resident-only): another configuration of BASELINE.json as the main workload; the benchmark's is C4 "
                         "and the others are reported under `configs`")
    ap.add_argument("--resident-only", action="store_true", help="tuning runs: the resident leg and the per-kernel times only, short JSON line")
    ap.add_argument("--lib", default=None, help="tuning runs (--resident-only): a kernel-variant build of the library")
    args = ap.parse_args()
    global MPDU_LEN, CONFIG, WORKLOAD
    MPDU_LEN = args.mpdu_len
    CONFIG = args.config
    WORKLOAD = CONFIGS[CONFIG]["workload"]
    if CONFIG != "C4" and not args.resident_only and args.impl == "ours":
        ap.error("--config other than C4 is for --resident-only tuning runs (the benchmark reports C2, C3, C5 under `configs`)")
    if args.impl == "reference":
        return main_reference(args)
    return main_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
