#!/usr/bin/env python
"""Resident throughput and latency of the fused path against batch size, per Viterbi kernel choice (tuning / evidence).

    python tools/batch_curve.py [--config C4] [--sizes 1,32,2048,...] [--modes auto,thread,warp]

One JSON line per (mode, size): Gbit/s with two slots in flight, per-stage CUDA-event times with one slot, and the
wall-clock latency of one w80211_rx_batch call (submit -> collect, device-resident input and output)."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import bench
    import w80211rx
    host = w80211rx.host
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C4")
    ap.add_argument("--sizes", default="1,32,256,2048,8192,14208,28416,56832")
    ap.add_argument("--modes", default="auto,thread,warp")
    ap.add_argument("--steps", type=int, default=6)
    a = ap.parse_args()
    bench.CONFIG = a.config
    dev = torch.device("cuda", 0)
    sizes = [int(x) for x in a.sizes.split(",")]
    NU = 128
    ub = bench.gen_unique(NU, 4)
    iq_u, pk_u, ch_u = host.pack_batch([ub])
    nant = len(iq_u)
    nsamp = int(pk_u["nsamp"][0])
    mlen = int(ub["mpdu"].shape[1]) if hasattr(ub["mpdu"], "shape") else len(ub["mpdu"][0])
    Pmax = max(sizes)
    reps = (Pmax + NU - 1) // NU
    iq_dev = [torch.from_numpy(np.tile(x, reps)[:Pmax * nsamp].copy()).to(dev) for x in iq_u]
    pk = np.zeros(Pmax, host.PKT_IN_DTYPE)
    pk["iq_offset"] = np.arange(Pmax, dtype=np.uint64) * nsamp
    pk["nsamp"] = nsamp
    pk["l_mcs"] = pk_u["l_mcs"][0]
    pk["l_len"] = pk_u["l_len"][0]
    pk_dev = torch.from_numpy(pk.view(np.uint8).copy()).to(dev)
    ch_dev = torch.from_numpy(np.tile(ch_u, (reps, 1))[:Pmax].view(np.float32).reshape(-1).copy()).to(dev)
    bytes_cap = Pmax * (mlen + 80) + 4096
    for mode in a.modes.split(","):
        rx = host.Receiver(0, nslots=2, viterbi=mode)
        L, h = rx.L, rx.h
        outs = []
        for _ in range(2):
            t = {"bytes": torch.empty(bytes_cap, dtype=torch.uint8, device=dev)}
            bo = host.BatchOut()
            bo.mem = host.MEM_DEVICE
            bo.bytes, bo.bytes_capacity = t["bytes"].data_ptr(), bytes_cap
            outs.append((bo, t))
        for P in sizes:
            bi = host.Receiver.make_in([x.data_ptr() for x in iq_dev], pk_dev.data_ptr(), ch_dev.data_ptr(), mem=host.MEM_DEVICE,
                                       nsamples=P * nsamp, npkt=P)

            def run(n, depth):
                bits = 0
                pend = []
                for i in range(n):
                    s = i % depth
                    if len(pend) == depth:
                        ps = pend.pop(0)
                        rx._check(L.w80211_rx_collect(h, ps))
                        bits += outs[ps][0].psdu_bits_ok
                    rx._check(L.w80211_rx_submit(h, s, C.byref(bi), C.byref(outs[s][0])))
                    pend.append(s)
                for ps in pend:
                    rx._check(L.w80211_rx_collect(h, ps))
                    bits += outs[ps][0].psdu_bits_ok
                return bits

            run(3, 2)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = a.steps if P >= 2048 else a.steps * 8
            e0.record()
            bits = run(n, 2)
            torch.cuda.synchronize()
            e1.record()
            e1.synchronize()
            ms = e0.elapsed_time(e1)
            # latency of one call, one slot
            lat = []
            st = {k: 0.0 for k in ("header", "demod", "viterbi", "traceback", "post", "total", "host_plan")}
            for _ in range(5):
                t0 = time.perf_counter()
                rx._check(L.w80211_rx_submit(h, 0, C.byref(bi), C.byref(outs[0][0])))
                rx._check(L.w80211_rx_collect(h, 0))
                lat.append((time.perf_counter() - t0) * 1e3)
                t = rx.timing(0)
                for k in st:
                    st[k] += getattr(t, k) / 5
            print(json.dumps({"config": a.config, "mode": mode, "packets": P, "gbit_s": bits / (ms * 1e-3) / 1e9, "ms_per_batch": ms / n,
                              "fcs_ok_per_batch": outs[0][0].n_crc_ok, "latency_ms_one_call": float(np.median(lat)),
                              "stage_ms": {k: round(v, 4) for k, v in st.items()}}), flush=True)
        rx.close()


if __name__ == "__main__":
    main()
