"""Parity of the CUDA path (through the C ABI) with the reference's own blocks on the same samples."""
import numpy as np
import pytest

import w80211rx
from util import compare_with_reference, ref_chain

gp = w80211rx.gen_ppdu
g = w80211rx.gen
host = w80211rx.host

pytestmark = pytest.mark.gpu

SISO_CASES = (
    [(g.F_L, m, 1, 60 + 37 * m, False) for m in range(8)]
    + [(g.F_HT, m, 1, 100 + 51 * m, False) for m in range(8)]
    + [(g.F_HT, m, 1, 333 + m, True) for m in (0, 3, 7)]
    + [(g.F_VHT, m, 1, 96 + 44 * m, False) for m in range(9)]
    + [(g.F_VHT, 8, 1, 500, True)]
)


@pytest.mark.parametrize("fmt,mcs,nss,length,sgi", SISO_CASES)
def test_siso_all_mcs(refgr, rx, fmt, mcs, nss, length, sgi):
    b = gp.gen_batch(fmt, mcs, nss, length, 5, snr_db=28, short_gi=sgi, seed=100 + mcs)
    res, msgs = compare_with_reference(refgr, rx, [b])
    assert len(msgs) == 5


def test_configs_c2_c3_c4(refgr, rx):
    """BASELINE.json configs 2-4 at their real packet sizes, a few packets each, one mixed batch."""
    bs = [gp.gen_batch(g.F_L, 7, 1, 1500, 3, snr_db=30, seed=2),
          gp.gen_batch(g.F_HT, 7, 1, 4095, 3, snr_db=30, short_gi=True, seed=3),
          gp.gen_batch(g.F_VHT, 8, 1, 4088, 3, snr_db=30, seed=4)]
    res, msgs = compare_with_reference(refgr, rx, bs)
    assert len(msgs) == 9
    for m, b in zip(msgs, [x for bb in bs for x in bb["mpdu"]]):
        assert m[3:-1] == b.tobytes()
