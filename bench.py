#!/usr/bin/env python
"""bench.py -- decoded PSDU throughput of the demod + decode hot path on B200.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[3], the configuration the 50 Gbit/s target is quoted on):
VHT MCS8 (256-QAM r3/4), 1 spatial stream, long GI, APEP 4092 (one 4088-byte MPDU per A-MPDU),
AWGN at 30 dB, synthetic.  A step is one pass of the whole path (header, demod, Viterbi, traceback,
descramble + FCS) over one batch of packets.  `value` is measured with the batch already in HBM;
`e2e` goes through the C ABI with host buffers (pinned), copies inside the timed region.
One JSON line on stdout (rank 0).
"""
import argparse
import ctypes as C
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
WORKLOAD = "C4: VHT MCS8 256-QAM r3/4 1SS long-GI, APEP 4092 B (MPDU 4088 B), AWGN 30 dB, 20 MHz"
MPDU_LEN = 4088
N_UNIQUE = 512


CONFIG = "C4"  # --config: the other single-antenna configurations of BASELINE.json, for tuning runs only


def gen_unique(n, seed):
    import w80211rx
    gp, g = w80211rx.gen_ppdu, w80211rx.gen
    if CONFIG == "C2":  # legacy 54 Mb/s, 1500 B
        return gp.gen_batch(g.F_L, 7, 1, 1500, n, snr_db=30, seed=seed)
    if CONFIG == "C3":  # HT MCS7 short GI, 4095 B (rate 5/6, trellis 32782)
        return gp.gen_batch(g.F_HT, 7, 1, 4095, n, snr_db=30, short_gi=True, seed=seed)
    return gp.gen_batch(g.F_VHT, 8, 1, MPDU_LEN, n, snr_db=30, seed=seed)


# ------------------------------------------------------------------ reference arm
def _ref_worker(args):
    """One oracle instance (the reference's blocks compiled unchanged) over its own packets."""
    seed, npkt = args
    from oracle import refgr
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import ref_chain
    b = gen_unique(npkt, seed)
    t0 = time.perf_counter()
    _, msgs = ref_chain(refgr, [b])
    dt = time.perf_counter() - t0
    sent = {x.tobytes() for x in b["mpdu"]}  # a packet whose FCS fails publishes nothing: match by content, not by position
    ok = sum(1 for m in msgs if m[3:-1] in sent)
    return dt, ok, npkt


def _port_worker(args):
    """Fallback when oracle/_ref is missing: the plain-C restatement of the decode block (oracle/port) over LLRs made
    from the same C4 packets' coded bits (+1/-1 with Gaussian noise).  The demod half (< 1 % of the CPU path) is not in it."""
    seed, npkt = args
    import w80211rx
    from oracle import port
    g = w80211rx.gen
    rng = np.random.default_rng(seed)
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
    return time.perf_counter() - t0, ok, npkt


def reference_kind():
    """'reference': the reference's own blocks (oracle/_ref); 'port': the plain-C restatement of the decode block; None."""
    from oracle import port, refgr
    return "reference" if refgr.available() else ("port" if port.available() else None)


def run_reference_sample(cores, pkts_per_core, seed0=1000):
    """-> (Gbit/s over all cores, packets ok, packets, wall seconds)."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        t0 = time.perf_counter()
        worker = _ref_worker if reference_kind() == "reference" else _port_worker
        res = pool.map(worker, [(seed0 + i, pkts_per_core) for i in range(cores)])
        wall = time.perf_counter() - t0
    ok = sum(r[1] for r in res)
    n = sum(r[2] for r in res)
    slowest = max(r[0] for r in res)  # decode time only (generation excluded), cores run side by side
    return ok * MPDU_LEN * 8 / slowest / 1e9, ok, n, slowest, wall


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
    per_core = args.ref_packets
    for _ in range(max(args.warmup, 0)):
        run_reference_sample(cores, 1)
    vals, tot_ok, tot_n, tsum = [], 0, 0, 0.0
    for k in range(args.steps):
        gbps, ok, n, slowest, _ = run_reference_sample(cores, per_core, 2000 + 100 * k)
        vals.append(gbps)
        tot_ok += ok
        tot_n += n
        tsum += slowest
    value = tot_ok * MPDU_LEN * 8 / tsum / 1e9
    what = "reference blocks (demod + decode) built unchanged, mock scheduler" if kind == "reference" else \
        "plain-C restatement of the decode block only (oracle/port; oracle/_ref missing), LLRs from the packets' coded bits"
    sample = "%d cores x %d packets per step, %d steps; %s" % (cores, per_core, args.steps, what)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tsum / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "packets_per_step": cores * per_core},
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


# ------------------------------------------------------------------ our arm
def main_gpu(args):
    import torch
    import w80211rx
    host = w80211rx.host
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    nsm = torch.cuda.get_device_properties(local).multi_processor_count
    # two full waves of the Viterbi kernel: 12 one-warp blocks per SM, one packet per lane
    P = args.packets if args.packets > 0 else nsm * 12 * 32 * 2
    PE = min(P, args.e2e_packets if args.e2e_packets > 0 else nsm * 12 * 32)  # host-fed legs: one full Viterbi wave per batch
    # ---- synthetic batch: N_UNIQUE distinct packets (own seed per rank), tiled to P packets per step
    ub = gen_unique(N_UNIQUE, 4 + 17 * rank)
    iq_u, pk_u, ch_u = host.pack_batch([ub])
    nsamp = int(pk_u["nsamp"][0])
    reps = (P + N_UNIQUE - 1) // N_UNIQUE
    pk = np.zeros(P, host.PKT_IN_DTYPE)
    pk["iq_offset"] = np.arange(P, dtype=np.uint64) * nsamp
    pk["nsamp"] = nsamp
    pk["l_mcs"] = pk_u["l_mcs"][0]
    pk["l_len"] = pk_u["l_len"][0]
    chan = np.ascontiguousarray(np.tile(ch_u, (reps, 1))[:P])
    # host copy (pinned) for the end-to-end leg, device copy for the resident leg
    iq_dev = torch.empty(P * nsamp, dtype=torch.complex64, device=dev)
    src = torch.from_numpy(iq_u[0]).to(dev)
    for r in range(reps):
        n = min(N_UNIQUE, P - r * N_UNIQUE) * nsamp
        iq_dev[r * N_UNIQUE * nsamp: r * N_UNIQUE * nsamp + n] = src[:n]
    pk_dev = torch.from_numpy(pk.view(np.uint8).copy()).to(dev)
    ch_dev = torch.from_numpy(chan.view(np.float32).reshape(-1).copy()).to(dev)
    iq_host = iq_dev[:PE * nsamp].cpu().pin_memory()
    pk_host = torch.from_numpy(pk[:PE].view(np.uint8).copy()).pin_memory()
    ch_host = torch.from_numpy(chan[:PE].view(np.float32).reshape(-1).copy()).pin_memory()
    # the same packets in the sc16 wire format: int16 pairs + one scale per packet
    q_u, pk16_u, _ = host.quantize_sc16(iq_u, pk_u)
    reps_e = (PE + N_UNIQUE - 1) // N_UNIQUE
    iq16_host = torch.from_numpy(np.tile(q_u[0], (reps_e, 1))[:PE * nsamp].copy()).pin_memory()
    pk16 = pk[:PE].copy()
    pk16["iq_scale"] = np.tile(pk16_u["iq_scale"], reps_e)[:PE]
    pk16_host = torch.from_numpy(pk16.view(np.uint8).copy()).pin_memory()
    q8_u, pk8_u, _ = host.quantize_sc16(iq_u, pk_u, bits=8)
    iq8_host = torch.from_numpy(np.tile(q8_u[0], (reps_e, 1))[:PE * nsamp].copy()).pin_memory()
    pk8 = pk[:PE].copy()
    pk8["iq_scale"] = np.tile(pk8_u["iq_scale"], reps_e)[:PE]
    pk8_host = torch.from_numpy(pk8.view(np.uint8).copy()).pin_memory()

    E2E_DEPTH = args.e2e_depth
    rx = host.Receiver(local, nslots=max(2, E2E_DEPTH))
    bytes_cap = P * (MPDU_LEN + 72) + 4096
    mp_cap = 2 * P + 64

    def dev_out():
        t = {"desc": torch.empty(P * host.DESC_DTYPE.itemsize, dtype=torch.uint8, device=dev),
             "bytes": torch.empty(bytes_cap, dtype=torch.uint8, device=dev),
             "psdu_off": torch.empty(P, dtype=torch.int32, device=dev),
             "mpdus": torch.empty(mp_cap * 16, dtype=torch.uint8, device=dev)}
        bo = host.BatchOut()
        bo.mem = host.MEM_DEVICE
        bo.desc, bo.bytes, bo.bytes_capacity = t["desc"].data_ptr(), t["bytes"].data_ptr(), bytes_cap
        bo.psdu_off, bo.mpdus, bo.mpdu_capacity = t["psdu_off"].data_ptr(), t["mpdus"].data_ptr(), mp_cap
        return bo, t

    bytes_cap_e, mp_cap_e = PE * (MPDU_LEN + 72) + 4096, 2 * PE + 64

    def host_out():
        t = {"desc": torch.empty(PE * host.DESC_DTYPE.itemsize, dtype=torch.uint8).pin_memory(),
             "bytes": torch.empty(bytes_cap_e, dtype=torch.uint8).pin_memory(),
             "psdu_off": torch.empty(PE, dtype=torch.int32).pin_memory(),
             "mpdus": torch.empty(mp_cap_e * 16, dtype=torch.uint8).pin_memory()}
        bo = host.BatchOut()
        bo.mem = host.MEM_HOST
        bo.desc, bo.bytes, bo.bytes_capacity = t["desc"].data_ptr(), t["bytes"].data_ptr(), bytes_cap_e
        bo.psdu_off, bo.mpdus, bo.mpdu_capacity = t["psdu_off"].data_ptr(), t["mpdus"].data_ptr(), mp_cap_e
        return bo, t

    bi_dev = host.Receiver.make_in([iq_dev.data_ptr()], pk_dev.data_ptr(), ch_dev.data_ptr(), mem=host.MEM_DEVICE,
                                   nsamples=P * nsamp, npkt=P)
    bi_host = host.Receiver.make_in([iq_host.data_ptr()], pk_host.data_ptr(), ch_host.data_ptr(), mem=host.MEM_DEVICE,
                                    nsamples=PE * nsamp, npkt=PE)
    bi_host.mem = host.MEM_HOST
    bi_host16 = host.Receiver.make_in([iq16_host.data_ptr()], pk16_host.data_ptr(), ch_host.data_ptr(), mem=host.MEM_DEVICE,
                                      nsamples=PE * nsamp, npkt=PE, iq_format=host.IQ_SC16)
    bi_host16.mem = host.MEM_HOST
    bi_host8 = host.Receiver.make_in([iq8_host.data_ptr()], pk8_host.data_ptr(), ch_host.data_ptr(), mem=host.MEM_DEVICE,
                                     nsamples=PE * nsamp, npkt=PE, iq_format=host.IQ_SC8)
    bi_host8.mem = host.MEM_HOST
    bo_d = [dev_out() for _ in range(2)]
    bo_h = [host_out() for _ in range(max(2, E2E_DEPTH))]
    L, h = rx.L, rx.h

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    stage = {k: 0.0 for k in ("header", "demod", "viterbi", "traceback", "post", "total", "host_plan", "host_collect")}
    counters = {"launches": 0}

    def run_steps(nsteps, bi, bos, tally, depth=2):
        """nsteps batches through w80211_rx_submit / collect with two slots in flight: the host plans batch i+1
        (it needs the header kernel's descriptors) while the GPU decodes batch i."""
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
            if tally:
                t = rx.timing(ps)
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

    # ---- resident leg: inputs already in HBM, results left in HBM
    clocks = ClockSampler(local)
    clocks.start()
    run_steps(max(args.warmup, 3), bi_dev, bo_d, False)
    barrier()
    t_from = clocks.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    bits, npk_ok = run_steps(args.steps, bi_dev, bo_d, False)
    barrier()
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1)
    # per-kernel durations (CUDA events on the slot's own stream): the same steps once more with ONE slot in
    # flight, because with two streams the events of one slot also span the other slot's kernels
    KSEQ = min(args.steps, 5)
    run_steps(KSEQ, bi_dev, bo_d, True, depth=1)
    torch.cuda.synchronize()
    t_to = clocks.mark()
    launches = counters["launches"] * args.steps // KSEQ
    if args.resident_only:
        print(json.dumps({"config": CONFIG, "value": bits / (ms * 1e-3) / 1e9, "ms_per_step": ms / args.steps, "packets": P, "lib": os.environ.get("W80211_LIB"),
                          "fcs_ok_per_step": npk_ok / args.steps, "clocks": clocks.stop(t_from, t_to),
                          "stage_ms_per_step": {k: round(v / KSEQ, 4) for k, v in stage.items()}}))
        rx.close()
        if dist is not None:
            dist.destroy_process_group()
        return 0

    # ---- end-to-end legs: pinned host buffers -> C ABI -> pinned host results, two slots in flight
    def e2e_leg(bi):
        run_steps(E2E_DEPTH, bi, bo_h, False, depth=E2E_DEPTH)
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        nbits, _ = run_steps(args.steps, bi, bo_h, False, depth=E2E_DEPTH)
        barrier()
        f1.record()
        f1.synchronize()
        return nbits, f0.elapsed_time(f1)

    e2e32_bits, e2e32_ms = e2e_leg(bi_host)    # fc32 samples on the wire (8 B per sample)
    e2e8_bits, e2e8_ms = e2e_leg(bi_host8)     # sc8 samples on the wire (2 B per sample): reported beside the headline
    e2e_bits, e2e_ms = e2e_leg(bi_host16)      # sc16 samples on the wire (4 B per sample): the headline host-fed figure
    clk = clocks.stop(t_from, t_to)

    # ---- front end (SURVEY.md 8(f) N3 + N1) on resident raw samples: every packet's whole PPDU with its own carrier offset,
    # 320 empty samples between packets.  presiso -> detect (trigger + sync) -> signal (L-SIG check, CFO-corrected copy), all in
    # HBM; the packets it hands over are then demodulated and decoded once.
    P_full = P

    def signal_leg():
        ppdu = ub["ppdu"][0]
        P = min(P_full, nsm * 12 * 32 // 2)  # a quarter of the resident batch: these stages scale linearly, HBM is nearly full by now
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
        KS = min(args.steps, 5)
        torch.cuda.synchronize()
        # presiso
        for _ in range(2):
            rx._check(L.w80211_presiso_batch(h, raw_dev.data_ptr(), nraw, host.MEM_DEVICE, ac_dev.data_ptr(), cj_dev.data_ptr()))
        t0 = time.perf_counter()
        for _ in range(KS):
            rx._check(L.w80211_presiso_batch(h, raw_dev.data_ptr(), nraw, host.MEM_DEVICE, ac_dev.data_ptr(), cj_dev.data_ptr()))
        pre_ms = (time.perf_counter() - t0) * 1e3 / KS
        # trigger + sync
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
        true_pos = np.arange(P, dtype=np.int64) * stride + gap + int(ub["sync_idx"])
        found = sy[:nflag]
        near = np.abs(found["pos"].astype(np.int64)[np.searchsorted(found["pos"].astype(np.int64), true_pos - 4).clip(0, max(nflag - 1, 0))] - true_pos) <= 4 \
            if nflag else np.zeros(P, bool)
        # signal stage on the detected flags
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
        _, ok2 = run_steps(1, bi2, bo_d, False)
        copy_bytes = float(npk_sig) * (nsamp * 16.0 + 1024.0)
        return {"packets_per_step": P, "raw_samples_per_step": nraw,
                "presiso_ms_host_wall": pre_ms, "presiso_gbs": nraw * 20.0 / (pre_ms * 1e-3) / 1e9,
                "detect_ms_host_wall": det_ms, "triggers": ntrig, "sync_flags": nflag, "true_packets_flagged": int(near.sum()),
                "signal_ms_host_wall": wall, "lsig_kernel_ms": lsig / KS, "cfo_copy_kernel_ms": copy / KS,
                "cfo_copy_gbs": copy_bytes / (copy / KS * 1e-3) / 1e9 if copy > 0 else None,
                "accepted": npk_sig, "fcs_ok_after_demod_decode": int(ok2),
                "note": "w80211_presiso_batch + w80211_detect_batch + w80211_signal_batch, raw samples and results in HBM; not part of "
                        "`value` (the north-star path starts at demod's input)"}

    sig_stage = signal_leg() if world == 1 else None
    h2d = PE * nsamp * 4 + pk_host.numel() + ch_host.numel() * 4
    h2d32 = PE * nsamp * 8 + pk_host.numel() + ch_host.numel() * 4
    d2h = int(bo_h[0][0].n_bytes) + int(bo_h[0][0].n_mpdu) * 16 + 64 + PE * host.DESC_DTYPE.itemsize

    # ---- max over ranks, totals over ranks
    vals = torch.tensor([ms, e2e_ms, e2e32_ms, e2e8_ms], dtype=torch.float64, device=dev)
    sums = torch.tensor([float(bits), float(npk_ok), float(e2e_bits), float(launches), float(e2e32_bits), float(e2e8_bits)],
                        dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    ms, e2e_ms, e2e32_ms, e2e8_ms = vals.tolist()
    bits, npk_ok, e2e_bits, launches, e2e32_bits, e2e8_bits = sums.tolist()
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    K = args.steps
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
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s"
    # per-packet sizes from the descriptor the header stage produced (C4: trellis 32760, total 43680, nSym 105, nCBPS 416)
    d0 = np.frombuffer(bo_h[0][1]["desc"].numpy(), host.DESC_DTYPE)[0]
    trellis, total, nsym, ncbps, nsymsamp = (int(d0[k]) for k in ("trellis", "total", "nsym", "ncbps", "nsymsamp"))
    # dominant kernel: the Viterbi forward pass.  Algorithmic bytes per trellis step: LLRs read
    # (total*4/trellis = 5.33 B at rate 3/4) + one 64-bit survivor word written.
    steps_per_launch = float(trellis) * P
    vit_ms = stage["viterbi"] / KSEQ
    alg_bytes = steps_per_launch * (total * 4.0 / trellis + 8.0)
    achieved = alg_bytes / (vit_ms * 1e-3) / 1e9
    sm_clk = clk["sm_mhz"] or float(peaks.get("sm_max_mhz", 1965.0))
    acs_per_s = steps_per_launch * 64 / (vit_ms * 1e-3)
    acs_peak = nsm * 128 * sm_clk * 1e6 / 3.0
    # per ACS the rate-3/4 body issues 1633 / 384 = 4.25 warp-instructions (834 FADD, 384 FMNMX, 384 SHF per six steps and
    # 64 states, cuobjdump -sass); one scheduler issues one warp-instruction per clock, so instruction issue bounds the kernel
    # before any pipe does (profiles/r1_pipebench.txt: FMNMX 1 and SHF 2 cycles of the ALU pipe per warp, FADD on the FMA pipe)
    INSTR_PER_ACS = 1633.0 / 384.0
    issue_peak = nsm * 4 * 32 * sm_clk * 1e6 / INSTR_PER_ACS
    alu_peak = nsm * 4 * 32 * sm_clk * 1e6 / 3.0
    demod_ms = stage["demod"] / KSEQ
    demod_bytes = float(P) * nsym * (nsymsamp * 8 + ncbps * 4)
    tb_ms = stage["traceback"] / KSEQ
    tb_bytes = steps_per_launch * 8.0 + steps_per_launch / 8.0
    line = {
        "metric": METRIC, "value": bits / (ms * 1e-3) / 1e9, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD if MPDU_LEN == 4088 else "PROFILING VARIANT, MPDU %d B, not the benchmark workload" % MPDU_LEN,
                   "packets_per_step_per_gpu": P, "unique_packets": N_UNIQUE, "slots_in_flight": 2,
                   "l2": "inputs larger than L2 (%.1f GB of IQ per step)" % (P * nsamp * 8 / 1e9), "parallelism": "packets sharded by rank, no collective"},
        "packets_per_s": npk_ok / (ms * 1e-3), "fcs_ok_per_step": npk_ok / K / world,
        "e2e": {"value": e2e_bits / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms / K, "packets_per_step_per_gpu": PE,
                "iq_wire_format": "sc16 + one scale per packet (SURVEY.md 8(f) N2), expanded to fc32 on the GPU",
                "slots_in_flight": E2E_DEPTH,
                "note": "pinned host buffers through w80211_rx_submit/collect; PCIe-bound, so smaller batches than the resident leg and one more slot in flight to keep the copy engine busy while the host plans"},
        "e2e_fc32": {"value": e2e32_bits / (e2e32_ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": h2d32, "d2h_bytes_per_step": d2h,
                     "ms_per_step": e2e32_ms / K, "iq_wire_format": "fc32, the reference flowgraph's sample type"},
        "e2e_sc8": {"value": e2e8_bits / (e2e8_ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": PE * nsamp * 2 + pk_host.numel() + ch_host.numel() * 4,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e8_ms / K,
                    "iq_wire_format": "sc8 + one scale per packet (UHD otw_format=sc8): 8-bit quantisation of the 30 dB samples, only FCS-valid bits are counted"},
        "gpu_launches": int(launches),
        "clocks": clk,
        "stage_ms_per_step": {k: v / KSEQ for k, v in stage.items()},
        "stage_note": "per-kernel CUDA-event durations from %d steps with one slot in flight, run right after the timed region" % KSEQ,
        "roofline": {"bound": "hbm", "kernel": "viterbi_kernel", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "algorithmic_bytes": alg_bytes,
                     "traffic": (traffic_tab["viterbi_kernel"]["dram_bytes_per_packet"] * P) if MPDU_LEN == 4088 and "viterbi_kernel" in traffic_tab else None,
                     "traffic_source": traffic_tab.get("source"), "peak_source": peak_src,
                     "note": "the kernel is FP32/ALU issue bound, not HBM bound: see roofline_compute"},
        "roofline_compute": {"kernel": "viterbi_kernel", "bound": "fp32+alu issue", "achieved": acs_per_s / 1e12, "peak": acs_peak / 1e12,
                             "unit": "T ACS/s", "frac": acs_per_s / acs_peak,
                             "peak_def": "SMs x 128 lanes x SM clock under load / 3 lane-ops per ACS (SURVEY.md 8(d))", "sm_mhz": sm_clk,
                             "issue_peak": issue_peak / 1e12, "issue_frac": acs_per_s / issue_peak,
                             "issue_def": "SMs x 4 schedulers x 32 lanes x clock / 4.25 warp-instructions per ACS (1633 per 6-step body of 384 ACS: 834 FADD, 384 FMNMX, 384 SHF)",
                             "alu_pipe_peak": alu_peak / 1e12, "alu_pipe_frac": acs_per_s / alu_peak,
                             "alu_pipe_def": "per ACS one FMNMX (1 cycle) and one SHF (2 cycles) on the ALU pipe, measured by tools/pipebench.cu: SMs x 4 SMSPs x 32 lanes x clock / 3 cycles"},
        "roofline_demod": {"kernel": "demod_kernel", "bound": "hbm", "achieved": demod_bytes / (demod_ms * 1e-3) / 1e9, "peak": hbm_peak,
                           "unit": "GB/s", "frac": demod_bytes / (demod_ms * 1e-3) / 1e9 / hbm_peak},
        "signal_stage": sig_stage,
        "roofline_traceback": {"kernel": "traceback_kernel", "bound": "hbm", "achieved": tb_bytes / (tb_ms * 1e-3) / 1e9, "peak": hbm_peak,
                               "unit": "GB/s", "frac": tb_bytes / (tb_ms * 1e-3) / 1e9 / hbm_peak},
    }
    # ---- CPU baseline beside it (rank 0, N = 1 only): the reference's own blocks on the host cores
    if world == 1 and not args.no_cpu:
        kind = reference_kind()
        if kind is not None:
            cores = host_cores()
            npc = max(args.ref_packets, 1024)
            gbps, ok, n, slowest, wall = run_reference_sample(cores, npc)
            line["cpu_baseline"] = {"value": gbps, "unit": UNIT, "cores": cores, "kind": kind,
                                    "sample": "%d packets of the same workload per core on %d cores, %.1f s; %s; %d/%d FCS ok"
                                              % (npc, cores, slowest, "reference demod + decode blocks built unchanged (oracle/_ref), mock scheduler"
                                                 if kind == "reference" else "plain-C restatement of the decode block (oracle/port), LLRs from coded bits", ok, n)}
    print(json.dumps(line))
    rx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


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
    ap.add_argument("--mpdu-len", type=int, default=4088, help="profiling only: shorter packets; the reported workload is the default")
    ap.add_argument("--ref-packets", type=int, default=512, help="packets per core and step for the CPU reference")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--config", default="C4", choices=["C2", "C3", "C4"], help="tuning runs (with --resident-only): another single-antenna configuration of BASELINE.json; the benchmark workload is C4")
    ap.add_argument("--resident-only", action="store_true", help="tuning runs: the resident leg and the per-kernel times only, short JSON line")
    args = ap.parse_args()
    global MPDU_LEN, CONFIG
    MPDU_LEN = args.mpdu_len
    CONFIG = args.config
    if CONFIG != "C4" and not args.resident_only and args.impl == "ours":
        ap.error("--config other than C4 is for --resident-only tuning runs")
    if args.impl == "reference":
        return main_reference(args)
    return main_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
