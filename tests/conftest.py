import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def refgr():
    """The reference's own blocks, built unchanged into oracle/_ref (test infrastructure)."""
    from oracle import refgr as r
    if not r.available():
        pytest.skip("oracle/_ref/libref80211.so not built (make -C oracle ref needs /root/reference)")
    return r


@pytest.fixture(scope="session", params=["thread", "warp"])
def rx(request):
    """One w80211_ctx on cuda:0, through the C ABI.  Fails (does not skip) when the library is missing.
    Every test that decodes runs twice: with the one-thread-per-packet Viterbi kernels and with the one-warp-per-packet
    ones (w80211_cfg.flags); both must equal the reference bit for bit."""
    import w80211rx
    r = w80211rx.host.Receiver(0, viterbi=request.param)
    yield r
    r.close()
