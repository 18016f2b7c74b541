"""the side paths of the one-antenna demod block -- VHT MU PPDUs demodulated for user position
`mupos` (demod_impl.cc:347-379, signalParserVhtA/B MU branches) and two-stream VHT NDPs whose raw VHT-LTF samples
travel to decode in the tag `mu2x1chan` (demod_impl.cc:238-271,393-409; decode_impl.cc:100-121).  Same bar as the
main path: tags exact, LLRs bit-identical, published blobs identical."""
import os

import numpy as np
import pytest

import w80211rx
from util import TAG_KEYS

gp, g, host = w80211rx.gen_ppdu, w80211rx.gen, w80211rx.host
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OURS_PATH = os.path.join(ROOT, "oracle", "_ours", "libours80211.so")

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ours(refgr):
    assert os.path.exists(OURS_PATH), "oracle/_ours/libours80211.so missing: run __graft_entry__.build()"
    return refgr.load_harness(OURS_PATH)


def packets_of(batches):
    pk = []
    for b in batches:
        pk += gp.to_packets(b)
    for i, p in enumerate(pk):
        p["seq"] = i
        p["cfo"], p["snr"], p["rssi"] = 0.5 * i, 25.0, -45.0
    return pk


NDP_CHAN = np.array([[1.0, 0.4 + 0.3j]])


@pytest.mark.parametrize("pos", [0, 1])
def test_mu_packets_through_the_c_abi(refgr, rx, pos):
    bs = [gp.gen_batch(g.F_VHT, m, 1, 200 + 90 * m, 3, snr_db=30, seed=40 + m, mu_pos=pos, short_gi=(m == 7)) for m in (0, 3, 5, 7, 8)]
    bs.append(gp.gen_batch(g.F_VHT, 4, 1, 400, 2, snr_db=30, seed=50))  # SU packets in the same batch are untouched by mu_pos
    iq, pkts, chan = host.pack_batch(bs)
    res = rx.rx(iq, pkts, chan, want_llr=True, mu_pos=pos)
    ref_llr = refgr.run_demod(refgr.make_sig_stream(packets_of(bs)), mupos=pos)
    ref_pkts = refgr.split_llr_stream(ref_llr)
    ref_msgs = refgr.run_decode(ref_llr)
    assert len(ref_pkts) == len(pkts) and all(res.desc["status"] == host.ST_OK)
    for p, r in enumerate(ref_pkts):
        for k in TAG_KEYS:
            assert int(res.desc[p][k]) == int(r[k]), (p, k)
        assert np.array_equal(res.packet_llr(p), r["llr"]), "packet %d: LLRs not bit-identical" % p
        assert abs(float(res.desc[p]["sssnr0"]) - float(r["sssnr0"])) <= 1e-3 * max(1.0, abs(float(r["sssnr0"])))
    assert res.messages() == ref_msgs and len(ref_msgs) == len(pkts)
    # the other user's position gives a wrong channel: nothing decodes, on either side
    res_w = rx.rx(iq[:1], pkts[:3], chan[:3], mu_pos=1 - pos)
    assert res_w.messages() == []


def test_ndp_descriptor(refgr, rx):
    b = gp.gen_batch(g.F_VHT, 0, 2, 16, 3, snr_db=30, seed=60, nrx=1, chan=NDP_CHAN, ndp=True)
    iq, pkts, chan = host.pack_batch([b])
    res = rx.demod(iq, pkts, chan)
    ref = refgr.run_demod(refgr.make_sig_stream(packets_of([b]))).tags()
    assert len(ref) == 3
    for p, (_, t) in enumerate(ref):
        d = res.desc[p]
        assert d["status"] == host.ST_NDP and d["nss"] == 2
        for k in ("format", "mcs", "len", "cr", "ampdu", "trellis"):
            assert int(d[k]) == int(t[k]), k
        assert abs(float(d["sssnr0"]) - float(t["sssnr0"])) <= 1e-3 * max(1.0, abs(float(t["sssnr0"])))
    # two streams of data on one antenna are refused, an NDP on the two-antenna path is dropped like the reference does
    b2 = gp.gen_batch(g.F_VHT, 3, 2, 300, 2, snr_db=30, seed=61, nrx=1, chan=NDP_CHAN)
    iq, pkts, chan = host.pack_batch([b2])
    assert all(rx.demod(iq, pkts, chan).desc["status"] == host.ST_UNSUPPORTED)
    b3 = gp.gen_batch(g.F_VHT, 0, 2, 16, 2, snr_db=30, seed=62, ndp=True)
    iq, pkts, chan = host.pack_batch([b3])
    assert all(rx.demod(iq, pkts, chan).desc["status"] == host.ST_DROP_GATE)
    assert len(refgr.run_demod(refgr.make_sig_stream(packets_of([b3]), nant=2)).tags()) == 0


@pytest.mark.parametrize("chunks", [(0, 8192), (777, 600)])
def test_blocks_ndp_and_mu_streams_equal_the_reference(refgr, ours, chunks):
    """demod(mupos=1) block: MU packets, SU packets and sounding packets interleaved; then decode on each side."""
    bs = [gp.gen_batch(g.F_VHT, 6, 1, 500, 2, snr_db=30, seed=70, mu_pos=1),
          gp.gen_batch(g.F_VHT, 0, 2, 16, 2, snr_db=30, seed=71, nrx=1, chan=NDP_CHAN, ndp=True),
          gp.gen_batch(g.F_HT, 4, 1, 300, 2, snr_db=30, seed=72),
          gp.gen_batch(g.F_VHT, 2, 1, 250, 2, snr_db=30, seed=73, mu_pos=1),
          gp.gen_batch(g.F_VHT, 0, 2, 16, 1, snr_db=30, seed=74, nrx=1, chan=NDP_CHAN, ndp=True)]
    pk = packets_of(bs)
    ref_out = refgr.run_demod(refgr.make_sig_stream(pk), mupos=1)
    our_out = ours.run_demod(ours.make_sig_stream(pk), mupos=1, chunk_in=chunks[0], chunk_out=chunks[1])
    tr, to = ref_out.tags(), our_out.tags()
    assert [o for o, _ in tr] == [o for o, _ in to] and len(tr) == 9
    dr, do = ref_out.data(), our_out.data()
    assert dr.shape == do.shape
    nndp = 0
    for (off, x), (_, y) in zip(to, tr):
        assert set(x) == set(y), set(x) ^ set(y)
        for k in y:
            if k == "mu2x1chan":
                assert np.array_equal(x[k], y[k])
                nndp += 1
            elif isinstance(y[k], float):
                assert abs(x[k] - y[k]) <= 1e-3 * max(1.0, abs(y[k])), k
            else:
                assert x[k] == y[k], k
        if "mu2x1chan" not in y:  # the 1024 items of a sounding packet are unwritten in the reference
            assert np.array_equal(do[off:off + y["total"]], dr[off:off + y["total"]])
    assert nndp == 3
    ref_msgs = refgr.run_decode(ref_out)
    our_msgs = ours.run_decode(our_out)
    assert our_msgs == ref_msgs
    assert sorted(len(m) for m in ref_msgs)[-3:] == [1027, 1027, 1027] and len(ref_msgs) == 9
