"""Parity of the detection stage (w80211_detect_batch) with the reference's trigger and sync blocks,
both fed the same presiso streams (tests/util.py:presiso restates examples/presiso.grc, a hierarchy of stock GNU Radio
blocks for which the reference holds no source): which samples get a sync flag (exact), rssi (exact), rad and snr (they
go through atan2f / cosf / sinf / log10: compared to a few ulp), and the whole chain from raw samples to published frames."""
import numpy as np
import pytest

import w80211rx
from util import build_stream, presiso

gp, g, host = w80211rx.gen_ppdu, w80211rx.gen, w80211rx.host

pytestmark = pytest.mark.gpu


def ref_detect(refgr, ac, cj, x):
    tr = refgr.run_trigger(ac)
    sy = refgr.run_sync(tr, cj, x)
    return [(o, t) for o, t in sy.tags()], tr, sy


def same_flags(mine, ref):
    assert len(mine) == len(ref), "sync flags: %d vs reference %d" % (len(mine), len(ref))
    for s, (o, t) in zip(mine, ref):
        assert int(s["pos"]) == o
        assert np.float32(s["rssi"]) == np.float32(t["rssi"])
        assert abs(float(s["rad"]) - t["rad"]) <= 1e-5 * abs(t["rad"]) + 2e-8, (float(s["rad"]), t["rad"])
        assert abs(float(s["snr"]) - t["snr"]) <= 1e-4 * max(1.0, abs(t["snr"]))


def stream_of(seed, noise=0.02, gap=600):
    bs = [gp.gen_batch(g.F_HT, 5, 1, 300, 4, snr_db=25, seed=seed), gp.gen_batch(g.F_L, 3, 1, 100, 3, snr_db=25, seed=seed + 1),
          gp.gen_batch(g.F_VHT, 4, 1, 500, 3, snr_db=25, seed=seed + 2)]
    raw, pos, rad = build_stream(bs, [0.01, -0.02, 0.004, -0.0007], gap=gap, noise=noise, seed=seed)
    return bs, raw, pos, rad


def test_detect_flags_equal_the_reference(refgr, rx):
    bs, raw, pos, rad = stream_of(1)
    ac, cj = presiso(raw[0])
    ref, _, _ = ref_detect(refgr, ac, cj, raw[0])
    mine = rx.detect(ac, cj, raw[0])
    same_flags(mine["syncs"], ref)
    assert len(ref) == len(pos) and np.all(np.abs(mine["syncs"]["pos"].astype(np.int64) - pos) <= 4)
    assert mine["next"] == len(ac) and mine["n_trigger"] >= len(pos)


def test_detect_strong_noise_and_close_packets(refgr, rx):
    """Noise 10 dB below the packets and packets only 250 samples apart: spurious runs above the threshold, triggers
    that land inside the 111 samples a sync pass consumes, countdowns that block the next run."""
    bs, raw, pos, rad = stream_of(7, noise=0.3, gap=250)
    ac, cj = presiso(raw[0])
    ref, _, _ = ref_detect(refgr, ac, cj, raw[0])
    mine = rx.detect(ac, cj, raw[0])
    same_flags(mine["syncs"], ref)
    assert len(ref) >= 8


def test_detect_to_frames(refgr, rx):
    """raw samples -> detect -> signal -> demod + decode on the GPU, against trigger -> sync -> signal -> demod -> decode"""
    bs, raw, pos, rad = stream_of(11)
    ac, cj = presiso(raw[0])
    _, _, sy = ref_detect(refgr, ac, cj, raw[0])
    blk = refgr.Block("signal")
    out = refgr.Stream(itemsize=8)
    blk.run([sy, refgr.Stream(raw[0])], [out])
    ref_llr = refgr.run_demod([out])
    ref_msgs = refgr.run_decode(ref_llr)
    det = rx.detect(ac, cj, raw[0])
    s = det["syncs"]
    sig = rx.signal(raw, s["pos"], rad=s["rad"], snr=s["snr"], rssi=s["rssi"])
    res = rx.rx(sig["iq"], sig["pkts"], sig["chan"])
    assert res.messages() == ref_msgs and len(ref_msgs) == 10


@pytest.mark.parametrize("step", [997, 2500, 4096])
def test_detect_streaming_equals_one_call(rx, step):
    """Buffers of `step` new samples, each call resuming at the previous call's `next` with its state: the flags are
    those of one call over the whole stream, wherever the cuts fall (inside runs, countdowns or sync windows)."""
    bs, raw, pos, rad = stream_of(21, noise=0.1, gap=300)
    x = raw[0]
    ac, cj = presiso(x)
    whole = rx.detect(ac, cj, x)["syncs"]
    got = []
    base, end, state = 0, 0, None
    while end < len(x):
        end = min(len(x), end + step)
        r = rx.detect(ac[base:end], cj[base:end], x[base:end], state=state, final=(end == len(x)))
        for s in r["syncs"]:
            got.append((int(s["pos"]) + base, float(s["rad"]), float(s["snr"]), float(s["rssi"])))
        state = r["state"]
        base += r["next"]
    assert got == [(int(s["pos"]), float(s["rad"]), float(s["snr"]), float(s["rssi"])) for s in whole]
    assert len(got) == len(pos)


def test_detect_empty_silence_and_bad_arguments(rx):
    r = rx.detect(np.zeros(0, np.float32), np.zeros(0, np.complex64), np.zeros(0, np.complex64))
    assert len(r["syncs"]) == 0 and r["next"] == 0
    z = np.zeros(5000, np.complex64)
    ac, cj = presiso(z)  # 0/0: NaN everywhere, never above the threshold
    r = rx.detect(ac, cj, z)
    assert len(r["syncs"]) == 0 and r["next"] == 5000 and r["state"].n_plateau == 0
    one = np.ones(5000, np.float32)  # a plateau that never ends: one trigger, no LTF behind it
    r = rx.detect(one, z, z)
    assert len(r["syncs"]) == 0 and r["n_trigger"] == 1 and r["state"].n_plateau == 5000 and r["state"].f_plateau_end == 1
    bad = host.DetectState()
    bad.count_down = 500
    with pytest.raises(host.W80211Error):
        rx.detect(one, z, z, state=bad)


def test_presiso_on_the_gpu(refgr, rx):
    """The GPU's presiso streams against the binary64 evaluation (1e-4 of the local scale), and the flags they lead to
    against the reference blocks fed the numpy streams: same packets, positions within one sample."""
    bs, raw, pos, rad = stream_of(31)
    x = raw[0]
    ac64, cj64 = presiso(x)
    ac, cj = rx.presiso(x)
    ok = np.isfinite(ac64) & (np.arange(len(x)) >= 80)
    assert np.all(np.abs(cj[ok] - cj64[ok]) <= 1e-4 * np.maximum(np.abs(cj64[ok]), 1e-3 * np.abs(cj64).max()))
    assert np.all(np.abs(ac[ok] - ac64[ok]) <= 1e-4 * np.maximum(ac64[ok], 1e-2))
    ref, _, _ = ref_detect(refgr, ac64, cj64, x)
    mine = rx.detect(ac, cj, x)["syncs"]
    assert len(mine) == len(ref) == len(pos)
    for s, (o, t) in zip(mine, ref):
        assert abs(int(s["pos"]) - o) <= 1 and abs(float(s["rad"]) - t["rad"]) <= 1e-4


@pytest.mark.parametrize("chunk", [0, 4000, 1537])
def test_presiso_against_gnuradios_running_sums(refgr, rx, chunk):
    """The stock GNU Radio blocks behind presiso.grc keep sequential binary32 running sums (carried over the whole stream in
    3.10, re-made at every work() call of at most max_iter = 4000 outputs in 3.8: oracle/port/presiso_port.c).  The GPU stage
    stays within 1e-4 of either, and the reference's trigger + sync blocks take the same decisions on either side's streams:
    the same flags at the same samples with the same rssi, rad and snr to the tolerance of same_flags."""
    from oracle import port
    bs, raw, pos, rad = stream_of(41, noise=0.05, gap=450)
    x = raw[0]
    ac_gr, cj_gr = port.presiso(x, chunk)
    ac, cj = rx.presiso(x)
    ok = np.isfinite(ac_gr) & (np.arange(len(x)) >= 80)
    # GNU Radio's carried binary32 sums drift by a few ulp of the largest value they have held: the floor of the comparison
    assert np.all(np.abs(cj[ok] - cj_gr[ok]) <= 1e-4 * np.abs(cj_gr[ok]) + 2e-6 * np.abs(cj_gr).max())
    busy = ok & (np.abs(cj_gr) > 1e-2 * np.abs(cj_gr).max())
    assert np.all(np.abs(ac[busy] - ac_gr[busy]) <= 1e-3 * np.maximum(ac_gr[busy], 1e-2))
    ref, _, _ = ref_detect(refgr, ac_gr, cj_gr, x)
    mine = rx.detect(ac, cj, x)["syncs"]
    assert len(mine) == len(ref) == len(pos)
    for s, (o, t) in zip(mine, ref):
        assert int(s["pos"]) == o
        assert abs(float(s["rad"]) - t["rad"]) <= 1e-4 and abs(float(s["snr"]) - t["snr"]) <= 1e-2 * max(1.0, abs(t["snr"]))
