"""The plain-C restatement of the reference's decode block (oracle/port/decode_port.c) against the reference's golden
fixture and against the reference blocks themselves (oracle/_ref).  CPU only."""
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
