"""Finds, with the reference's own blocks (oracle/_ref, CPU), the SNR at which each MCS sits in its waterfall -- the
packet-delivery ratio of tools/performance/perf_siso.py:87-157 / perf_sumimo.py:89-123 (100 packets of ~570 bytes per
MCS over an SNR sweep) is neither 0 nor 1 there.  tests/test_gpu_pdr.py runs its sweep at these SNRs - 1, + 0, + 1 dB.
Usage: python tools/pdr_calibrate.py  -> prints the table that test holds (needs /root/reference built into oracle/_ref)."""
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
LENGTH = 570
CASES = [("L", m, 1) for m in range(8)] + [("HT", m, 1) for m in range(8)] + [("VHT", m, 1) for m in range(9)] + \
        [("HT", m, 2) for m in range(8, 16)] + [("VHT", m, 2) for m in range(9)]


def pdr(args):
    fmt, mcs, nss, snr, n, seed = args
    import w80211rx
    from oracle import refgr
    gp, g = w80211rx.gen_ppdu, w80211rx.gen
    f = {"L": g.F_L, "HT": g.F_HT, "VHT": g.F_VHT}[fmt]
    b = gp.gen_batch(f, mcs, nss, LENGTH, n, snr_db=snr, seed=seed)
    pk = gp.to_packets(b)
    msgs = refgr.run_decode(refgr.run_demod(refgr.make_sig_stream(pk, nant=b["nrx"])))
    return len(msgs) / n


def main():
    n = 24
    with mp.get_context("fork").Pool(8) as pool:
        out = {}
        for fmt, mcs, nss in CASES:
            lo, hi = 0.0, 40.0
            snrs = np.arange(lo, hi + 0.1, 2.0)
            r = pool.map(pdr, [(fmt, mcs, nss, float(s), n, 1000 + i) for i, s in enumerate(snrs)])
            # first SNR with PDR >= 0.5, refined by 0.5 dB steps around it
            k = next((i for i, v in enumerate(r) if v >= 0.5), len(r) - 1)
            fine = np.arange(snrs[max(k - 1, 0)], snrs[k] + 0.1, 0.5)
            r2 = pool.map(pdr, [(fmt, mcs, nss, float(s), 2 * n, 2000 + i) for i, s in enumerate(fine)])
            k2 = int(np.argmin([abs(v - 0.5) for v in r2]))
            out[(fmt, mcs, nss)] = (float(fine[k2]), r2[k2])
            print(fmt, mcs, nss, "-> %.1f dB (PDR %.2f)" % out[(fmt, mcs, nss)], flush=True)
    print("WATERFALL_DB = {")
    for (fmt, mcs, nss), (s, v) in out.items():
        print('    ("%s", %d, %d): %.1f,' % (fmt, mcs, nss, s))
    print("}")


if __name__ == "__main__":
    main()
