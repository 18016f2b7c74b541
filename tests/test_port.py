"""The plain-C restatements of the reference's decode block (oracle/port/decode_port.c) and of its per-symbol
demodulation (oracle/port/demod_port.c) against the reference's golden fixture and against the reference blocks
themselves (oracle/_ref).  CPU only."""
import os

import numpy as np
import pytest

import w80211rx
from oracle import port

gp, g = w80211rx.gen_ppdu, w80211rx.gen
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
pytestmark = pytest.mark.skipif(not port.available(), reason="oracle/libport80211.so not built (make -C oracle port)")


@pytest.mark.parametrize("name", ["fixture_siso.npz", "fixture_mimo.npz"])
def test_port_decodes_the_reference_fixture(name):
    """LLR stream of tools/sig80211GenMultiple*.bin (committed fixture) -> the 94-byte frame of tools/debug.txt:5."""
    z = np.load(os.path.join(GOLD, name))
    fmt, mcs, ln, cr, ampdu, trellis, total = [int(x) for x in z["tags"]]
    msgs, nfail = port.decode_packet({"llr": z["llr"], "format": fmt, "mcs": mcs, "len": ln, "cr": cr, "ampdu": ampdu, "trellis": trellis})
    assert nfail == 0 and msgs == [z["msg"].tobytes()]
    assert msgs[0][3:-1] == np.load(os.path.join(GOLD, "frame94.npy")).tobytes()


def test_port_trellis_tables_are_the_reference_tables(refgr):
    nx, op = refgr.sv_tables()
    pn, pl = port.trellis_tables()
    assert np.array_equal(nx, pn) and np.array_equal(op, pl)


def test_port_matches_reference_decode_on_every_code_rate(refgr):
    """legacy / HT / VHT packets of all four code rates, a multi-subframe A-MPDU and wrong FCS: same messages as the
    reference's decode block; the Viterbi bits equal the reference's free-function decoder on rate 1/2."""
    bs = [gp.gen_batch(g.F_L, 0, 1, 120, 2, snr_db=12, seed=901),       # 1/2
          gp.gen_batch(g.F_L, 6, 1, 400, 2, snr_db=28, seed=902),       # 2/3
          gp.gen_batch(g.F_HT, 4, 1, 700, 2, snr_db=25, seed=903),      # 3/4
          gp.gen_batch(g.F_HT, 7, 1, 900, 2, snr_db=30, seed=904),      # 5/6
          gp.gen_batch(g.F_VHT, 8, 1, 1500, 2, snr_db=30, seed=905),
          gp.gen_batch(g.F_VHT, 4, 1, 300, 2, snr_db=30, seed=906, corrupt_fcs=True),
          gp.gen_batch(g.F_L, 3, 1, 200, 2, snr_db=3, seed=907)]        # noisy: most FCS fail, same verdicts
    pk = []
    for b in bs:
        pk += gp.to_packets(b)
    llr = refgr.run_demod(refgr.make_sig_stream(pk))
    want = refgr.run_decode(llr)
    got = port.run_decode(refgr.split_llr_stream(llr))
    assert got == want and len(want) >= 8
    x = np.random.default_rng(3).normal(0, 1, 2 * 500).astype(np.float32)
    bits, used = port.viterbi(x, 500, 0)
    assert used == 1000 and np.array_equal(bits, refgr.sv_decode(x, 500))


def test_demod_port_reproduces_the_reference_fixture():
    """Samples + legacy channel of tools/sig80211GenMultipleSiso_1x1_0.bin (committed fixture) -> the LLR stream the
    reference's demod block produced from them, bit for bit (legacy BPSK, 33 symbols)."""
    z = np.load(os.path.join(GOLD, "fixture_siso.npz"))
    fmt, mcs, ln, cr, ampdu, trellis, total = [int(x) for x in z["tags"]]
    assert fmt == 0 and total == 33 * 48
    llr = port.demod_packet(z["post"][0], z["chan"], 0, 1, total // 48)
    assert np.array_equal(llr, z["llr"])


NB_L = [1, 1, 2, 2, 4, 4, 6, 6]
NB_NL = [1, 2, 2, 4, 4, 6, 6, 6, 8, 8]


@pytest.mark.parametrize("fmt,mcs,length,sgi", [(0, 0, 60, False), (0, 3, 200, False), (0, 5, 333, False), (0, 7, 1500, False),
                                                (1, 0, 40, False), (1, 2, 300, False), (1, 4, 700, False), (1, 7, 1200, False),
                                                (1, 7, 900, True), (2, 1, 100, False), (2, 5, 1000, False), (2, 8, 2000, False),
                                                (2, 8, 1500, True)])
def test_demod_port_matches_reference_demod(refgr, fmt, mcs, length, sgi):
    """Generated packets of every modulation, long and short GI: the port's LLRs of the data symbols equal the
    reference demod block's (oracle/_ref), element for element."""
    from util import ref_chain
    f = [g.F_L, g.F_HT, g.F_VHT][fmt]
    b = gp.gen_batch(f, mcs, 1, length, 3, snr_db=25, short_gi=sgi, seed=900 + 17 * fmt + mcs)
    ref_pkts, _ = ref_chain(refgr, [b])
    nb = NB_L[mcs] if fmt == 0 else NB_NL[mcs]
    nsd = 48 if fmt == 0 else 52
    for p, r in enumerate(ref_pkts):
        assert int(r["format"]) == fmt
        nsym = int(r["total"]) // (nsd * nb)
        got = port.demod_packet(b["post"][0][p], b["chan"][p], fmt, nb, nsym, short_gi=sgi)
        assert np.array_equal(got, r["llr"][: len(got)]) and len(got) == int(r["total"])


def test_presiso_port_variants_agree_with_a_binary64_evaluation(refgr):
    """oracle/port/presiso_port.c (GNU Radio's binary32 running sums, carried or re-made per work() call) against the
    binary64 evaluation of tests/util.py, and the reference's trigger + sync blocks fed each: same flags (CPU only)."""
    import w80211rx
    from util import build_stream, presiso
    gp, g = w80211rx.gen_ppdu, w80211rx.gen
    bs = [gp.gen_batch(g.F_HT, 4, 1, 200, 3, snr_db=25, seed=51), gp.gen_batch(g.F_L, 2, 1, 90, 3, snr_db=25, seed=52)]
    raw, pos, rad = build_stream(bs, [0.01, -0.015], gap=500, noise=0.05, seed=5)
    x = raw[0]
    ac64, cj64 = presiso(x)
    flags = []
    for chunk in (0, 4000, 777):
        ac, cj = port.presiso(x, chunk)
        ok = np.isfinite(ac64) & (np.arange(len(x)) >= 80)
        # a carried binary32 sum drifts by a few ulp of the largest value it has held: the floor of the comparison
        assert np.all(np.abs(cj[ok] - cj64[ok]) <= 1e-4 * np.abs(cj64[ok]) + 2e-6 * np.abs(cj64).max())
        busy = ok & (np.abs(cj64) > 1e-2 * np.abs(cj64).max())
        assert np.all(np.abs(ac[busy] - ac64[busy]) <= 1e-3 * np.maximum(ac64[busy], 1e-2))
        sy = refgr.run_sync(refgr.run_trigger(ac), cj, x)
        flags.append([o for o, _ in sy.tags()])
    assert flags[0] == flags[1] == flags[2] and len(flags[0]) == len(pos)
