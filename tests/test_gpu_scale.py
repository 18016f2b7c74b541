"""Parity at the size the benchmark runs at: full-length packets of every BASELINE.json configuration (C2 .. C5) and of
the two code rates none of them uses, 72 per case -- two whole 32-packet Viterbi groups and a ragged third per code rate --
a third of them pushed into their waterfall so that some fail.  LLRs bit-identical to the reference demod's, messages
(hence the set of packets whose FCS holds) identical to the reference decode's (oracle/_ref, the reference blocks built
unchanged; about 30 ms of CPU per 4 KB packet)."""
import numpy as np
import pytest

import w80211rx
from util import compare_with_reference, ref_chain

gp, g, host = w80211rx.gen_ppdu, w80211rx.gen, w80211rx.host
pytestmark = pytest.mark.gpu

# name: (format, mcs, streams, MPDU bytes, short GI, code rate, SNR of the degraded third)
CASES = {
    "C2": (g.F_L, 7, 1, 1500, False, "3/4", 19.5),
    "C3": (g.F_HT, 7, 1, 4095, True, "5/6", 22.0),
    "C4": (g.F_VHT, 8, 1, 4088, False, "3/4", 26.5),
    "C5": (g.F_HT, 15, 2, 4095, False, "5/6", 25.0),
    "r12": (g.F_VHT, 3, 1, 4088, False, "1/2", 12.0),
    "r23": (g.F_VHT, 5, 1, 4000, False, "2/3", 19.0),
}
_REF = {}


def batches_of(name):
    fmt, mcs, nss, length, sgi, _, low = CASES[name]
    seed = 5000 + 10 * sorted(CASES).index(name)
    return [gp.gen_batch(fmt, mcs, nss, length, 48, snr_db=30, short_gi=sgi, seed=seed),
            gp.gen_batch(fmt, mcs, nss, length, 24, snr_db=low, short_gi=sgi, seed=seed + 1)]


@pytest.mark.parametrize("name", sorted(CASES))
def test_full_size_groups_equal_the_reference(refgr, rx, name):
    bs = batches_of(name)
    if name not in _REF:  # the rx fixture runs every test under both Viterbi layouts: the reference side runs once
        _REF[name] = ref_chain(refgr, bs)
    res, msgs = compare_with_reference(refgr, rx, bs, ref=_REF[name])
    sent = [m.tobytes() for b in bs for m in b["mpdu"]]
    assert all(m[3:-1] in set(sent) for m in msgs)
    assert [m[3:-1] for m in msgs[:48]] == sent[:48], "the 30 dB packets all decode"
    assert 48 <= len(msgs) <= 72
    # the decoder saw what the docstring promises: >= 2 full groups of this code rate
    ok = res.desc["status"] == host.ST_OK
    assert int(ok.sum()) >= 64 and len(set(res.desc["cr"][ok])) == 1


def test_the_degraded_thirds_do_fail_somewhere(refgr):
    """The sweep above is only a fail-SET test if something fails: over all cases the reference loses packets of the
    degraded thirds (not all of them), so FCS verdicts, not just clean decodes, are being compared."""
    if len(_REF) < len(CASES):
        pytest.skip("runs after the parity cases")
    lost = {n: 72 - len(_REF[n][1]) for n in CASES}
    assert sum(1 for v in lost.values() if 0 < v < 24) >= 3, lost
