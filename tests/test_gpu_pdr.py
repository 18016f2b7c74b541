"""The reference's own acceptance method, as a parity test: packet delivery per MCS over an SNR sweep
(tools/performance/perf_siso.py:87-157, perf_sumimo.py:89-123: 100 packets of a ~570-byte MPDU per MCS, legacy / HT / VHT,
one and two streams, counted from the decode block's 'crc32 correct' lines).  Here every MCS runs at three SNRs around its
waterfall (WATERFALL_DB: where the reference's PDR is about one half, found with the reference's blocks by
tools/pdr_calibrate.py), where delivery is neither 0 nor 1, and the set of published frames -- not just the count -- must
equal the reference's (oracle/_ref, the reference blocks built unchanged)."""
import numpy as np
import pytest

import w80211rx
from util import ref_chain

gp, g, host = w80211rx.gen_ppdu, w80211rx.gen, w80211rx.host
pytestmark = pytest.mark.gpu

LENGTH, NPKT = 570, 100
WATERFALL_DB = {
    ("L", 0, 1): 2.5, ("L", 1, 1): 4.0, ("L", 2, 1): 5.5, ("L", 3, 1): 7.0, ("L", 4, 1): 10.0, ("L", 5, 1): 13.0, ("L", 6, 1): 17.0,
    ("L", 7, 1): 18.5,
    ("HT", 0, 1): 4.5, ("HT", 1, 1): 7.0, ("HT", 2, 1): 9.0, ("HT", 3, 1): 11.5, ("HT", 4, 1): 14.5, ("HT", 5, 1): 18.5, ("HT", 6, 1): 20.0,
    ("HT", 7, 1): 21.0,
    ("VHT", 0, 1): 4.5, ("VHT", 1, 1): 7.0, ("VHT", 2, 1): 8.5, ("VHT", 3, 1): 11.0, ("VHT", 4, 1): 14.5, ("VHT", 5, 1): 18.0,
    ("VHT", 6, 1): 19.5, ("VHT", 7, 1): 21.0, ("VHT", 8, 1): 25.5,
    ("HT", 8, 2): 7.0, ("HT", 9, 2): 9.5, ("HT", 10, 2): 11.5, ("HT", 11, 2): 14.0, ("HT", 12, 2): 17.0, ("HT", 13, 2): 21.0,
    ("HT", 14, 2): 22.5, ("HT", 15, 2): 24.0,
    ("VHT", 0, 2): 7.5, ("VHT", 1, 2): 9.5, ("VHT", 2, 2): 11.5, ("VHT", 3, 2): 14.0, ("VHT", 4, 2): 17.5, ("VHT", 5, 2): 21.5,
    ("VHT", 6, 2): 22.5, ("VHT", 7, 2): 24.0, ("VHT", 8, 2): 28.0,
}
FMT = {"L": g.F_L, "HT": g.F_HT, "VHT": g.F_VHT}
_REF = {}
_PDR = {}


def sweep_batches(fmt, mcs, nss):
    w = WATERFALL_DB[(fmt, mcs, nss)]
    seed = 9000 + 100 * list(WATERFALL_DB).index((fmt, mcs, nss))
    return [gp.gen_batch(FMT[fmt], mcs, nss, LENGTH, NPKT, snr_db=w + d, seed=seed + i) for i, d in enumerate((-1.0, 0.0, 1.0))]


@pytest.mark.parametrize("fmt,mcs,nss", sorted(WATERFALL_DB))
def test_published_set_over_the_snr_sweep(refgr, rx, fmt, mcs, nss):
    bs = sweep_batches(fmt, mcs, nss)
    key = (fmt, mcs, nss)
    if key not in _REF:  # once for both Viterbi layouts of the rx fixture
        _REF[key] = [ref_chain(refgr, [b])[1] for b in bs]
    got_n = []
    for b, want in zip(bs, _REF[key]):
        iq, pkts, chan = host.pack_batch([b])
        res = rx.rx(iq, pkts, chan)
        got = res.messages()
        assert got == want, "published frames differ from the reference's (%d vs %d)" % (len(got), len(want))
        sent = {m.tobytes() for m in b["mpdu"]}
        assert all(m[3:-1] in sent for m in got)
        got_n.append(len(got))
    _PDR[key] = got_n
    assert got_n[0] <= got_n[2] + 10  # delivery grows with SNR (allowing for the spread of 100 packets)


def test_the_sweep_sits_in_the_waterfalls():
    """Almost every MCS must have lost some packets and delivered some at its middle SNR: otherwise the sweep compares
    only trivial sets."""
    if len(_PDR) < len(WATERFALL_DB):
        pytest.skip("runs after the sweep")
    inside = [k for k, v in _PDR.items() if 0 < v[1] < NPKT]
    assert len(inside) >= len(WATERFALL_DB) - 4, {k: v for k, v in _PDR.items() if k not in inside}
