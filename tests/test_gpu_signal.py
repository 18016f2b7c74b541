"""Parity of the signal stage (w80211_signal_batch) with the reference's signal / signal2
blocks on the same raw samples and sync flags: which flags become packets, the tags, the legacy channel, the
CFO-corrected samples, and -- through the rest of the chain -- the LLRs and the published messages.

The rotation factors are cosf/sinf of a binary32 angle.  A host libm and the GPU's sincosf both return them to
within 1 ulp but not always the same neighbour, so samples, channel and LLRs are compared to 1e-4 relative (the
tolerance of BASELINE.json), the integer tags and the decoded bytes exactly."""
import numpy as np
import pytest

import w80211rx
from util import llr_close, TAG_KEYS, build_stream

gp = w80211rx.gen_ppdu
g = w80211rx.gen
host = w80211rx.host

pytestmark = pytest.mark.gpu


def ref_signal(refgr, raw, pos, rad, snr=17.5, rssi=-42.0):
    """-> (list of packet dicts from the reference block's output stream: tags + 'iq' per antenna, the streams)"""
    outs = refgr.run_signal(raw, [int(x) for x in pos], rad=rad, snr=snr, rssi=rssi)
    data = [o.data() for o in outs]
    pk = []
    for off, t in outs[0].tags():
        d = dict(t)
        d["offset"] = off
        d["iq"] = [x[off:off + d["nsamp"]] for x in data]
        pk.append(d)
    return pk, outs


def close(a, b, tol=1e-4):
    a = np.asarray(a)
    b = np.asarray(b)
    scale = max(float(np.abs(b).max()), 1e-30) if b.size else 1.0
    return a.shape == b.shape and bool(np.all(np.abs(a - b) <= tol * scale))


def check_against_reference(refgr, rx, raw, pos, rad, expect=None):
    ref_pk, ref_streams = ref_signal(refgr, raw, pos, rad)
    mine = rx.signal(raw, pos, rad=rad, snr=17.5, rssi=-42.0)
    if expect is not None:
        assert len(ref_pk) == expect
    assert len(mine["pkts"]) == len(ref_pk), "accepted packets: %d vs reference %d" % (len(mine["pkts"]), len(ref_pk))
    for k, (pk, r) in enumerate(zip(mine["pkts"], ref_pk)):
        assert (int(pk["l_mcs"]), int(pk["l_len"]), int(pk["nsamp"])) == (r["mcs"], r["len"], r["nsamp"]), k
        assert np.float32(pk["cfo"]) == np.float32(r["cfo"]) and np.float32(pk["snr"]) == np.float32(r["snr"])
        assert np.float32(pk["rssi"]) == np.float32(r["rssi"])
        assert close(mine["chan"][k], r["chan"]), "packet %d: legacy channel differs" % k
        assert np.array_equal(mine["chan"][k] == 0, r["chan"] == 0)
        o = int(pk["iq_offset"])
        for a in range(len(raw)):
            assert close(mine["iq"][a][o:o + int(pk["nsamp"])], r["iq"][a]), "packet %d antenna %d: samples differ" % (k, a)
    return mine, ref_pk, ref_streams


def chain_equal(refgr, rx, mine, ref_streams):
    """demod + decode of both sides: tags exact, LLRs to 1e-4, messages identical"""
    res = rx.rx(mine["iq"], mine["pkts"], mine["chan"], want_llr=True)
    llr = refgr.run_demod(ref_streams)
    ref_pkts = refgr.split_llr_stream(llr)
    ref_msgs = refgr.run_decode(llr)
    ok = [p for p in range(len(mine["pkts"])) if res.desc[p]["status"] == host.ST_OK]
    assert len(ok) == len(ref_pkts)
    for p, r in zip(ok, ref_pkts):
        d = res.desc[p]
        for key in TAG_KEYS:
            assert int(d[key]) == int(r[key]), (p, key)
        scale = {1: 1.0, 2: 1.0, 4: 2.0, 6: 4.0, 8: 8.0}[int(d["nbpscs"])]
        assert llr_close(res.packet_llr(p), r["llr"], scale), "packet %d LLRs differ by more than 1e-4 relative" % p
    assert res.messages() == ref_msgs
    return ref_msgs


def test_signal_siso_mixed_formats_with_cfo(refgr, rx):
    bs = [gp.gen_batch(g.F_L, 4, 1, 300, 3, snr_db=30, seed=1),
          gp.gen_batch(g.F_HT, 5, 1, 700, 3, snr_db=30, seed=2),
          gp.gen_batch(g.F_VHT, 8, 1, 1200, 3, snr_db=32, seed=3),
          gp.gen_batch(g.F_L, 0, 1, 40, 2, snr_db=25, seed=4)]
    raw, pos, rad = build_stream(bs, [0.0, 0.011, -0.0207, 0.0031, -0.0005])
    mine, ref_pk, ref_streams = check_against_reference(refgr, rx, raw, pos, rad, expect=11)
    assert mine["next"] == len(raw[0]) and mine["nsync_used"] == len(pos) and mine["n_rejected"] == 0
    msgs = chain_equal(refgr, rx, mine, ref_streams)
    assert len(msgs) == 11


def test_signal2_two_antennas(refgr, rx):
    h = np.array([[1.0, 0.35 + 0.2j], [0.25 - 0.3j, 0.9]])
    bs = [gp.gen_batch(g.F_HT, 11, 2, 500, 3, snr_db=35, seed=5, chan=h),
          gp.gen_batch(g.F_VHT, 4, 2, 800, 3, snr_db=35, seed=6, chan=h)]
    raw, pos, rad = build_stream(bs, [0.004, -0.013])
    mine, ref_pk, ref_streams = check_against_reference(refgr, rx, raw, pos, rad, expect=6)
    msgs = chain_equal(refgr, rx, mine, ref_streams)
    assert len(msgs) == 6


def test_signal_false_flags_and_overlaps(refgr, rx):
    """Flags on noise fail the L-SIG check (the block resumes 80 samples later); flags inside an accepted packet
    and within 80 samples of a rejected one are never examined; a flag without tags is skipped."""
    bs = [gp.gen_batch(g.F_L, 6, 1, 400, 4, snr_db=30, seed=7)]
    raw, pos, rad = build_stream(bs, [0.002], gap=700, noise=0.05)
    extra = []
    for p in pos:
        extra += [p - 600, p - 560, p - 300, p + 100, p + 900]  # noise, inside the 80-sample skip, noise, inside the packet x2
    allpos = np.sort(np.concatenate([pos, np.array(extra, np.int64)]))
    rads = np.full(len(allpos), 0.002, np.float32)
    ref_pk, ref_streams = ref_signal(refgr, raw, allpos, rads)
    mine = rx.signal(raw, allpos, rad=rads, snr=17.5, rssi=-42.0)
    assert len(ref_pk) == len(mine["pkts"])
    for pk, r in zip(mine["pkts"], ref_pk):
        assert (int(pk["l_mcs"]), int(pk["l_len"]), int(pk["nsamp"])) == (r["mcs"], r["len"], r["nsamp"])
    # every true packet is found, and the accepted flags are the true ones unless noise passed the 1-bit parity check
    true_found = [int(allpos[i]) in set(int(x) for x in pos) for i in mine["sync_of_pkt"]]
    assert sum(true_found) == 4
    assert mine["n_rejected"] > 0
    # untagged flag: skipped by both
    tagged = np.ones(len(pos), bool)
    tagged[1] = False
    m2 = rx.signal(raw, pos, rad=0.002, tagged=tagged)
    assert len(m2["pkts"]) == 3 and list(m2["sync_of_pkt"]) == [0, 2, 3]


def test_signal_streaming_carry_over(rx):
    """A buffer that ends inside a packet: the packet is left for the next call (next = its flag), and the two
    calls together give exactly what one call over the whole stream gives."""
    bs = [gp.gen_batch(g.F_HT, 3, 1, 600, 4, snr_db=30, seed=8)]
    raw, pos, rad = build_stream(bs, [0.006, -0.004])
    whole = rx.signal(raw, pos, rad=rad)
    assert len(whole["pkts"]) == 4
    cut = int(pos[2]) + 500
    first = rx.signal([raw[0][:cut]], pos[pos + 224 <= cut], rad=rad[pos + 224 <= cut], final=False)
    assert len(first["pkts"]) == 2 and first["next"] == int(pos[2]) and first["nsync_used"] == 2
    nxt = first["next"]
    second = rx.signal([raw[0][nxt:]], pos[2:] - nxt, rad=rad[2:])
    assert len(second["pkts"]) == 2
    got = np.concatenate([first["iq"][0], second["iq"][0]])
    assert np.array_equal(got, whole["iq"][0])
    assert np.array_equal(np.concatenate([first["chan"], second["chan"]]), whole["chan"])
    # final=True drops the cut packet instead
    last = rx.signal([raw[0][:cut]], pos[:3], rad=rad[:3], final=True)
    assert len(last["pkts"]) == 2 and last["next"] == cut


def test_signal_empty_and_bad_arguments(rx):
    raw = [np.zeros(1000, np.complex64)]
    r = rx.signal(raw, np.zeros(0, np.int64))
    assert len(r["pkts"]) == 0 and r["next"] == 1000
    r = rx.signal(raw, [900])  # fewer than 224 samples after the flag, final: dropped
    assert len(r["pkts"]) == 0
    r = rx.signal(raw, [100, 400])  # zeros: 0/0 in the equaliser gives NaN LLRs, the check fails like the reference's
    assert len(r["pkts"]) == 0
    with pytest.raises(host.W80211Error):
        rx.signal(raw, [500, 100])
