"""CPU tests of the oracle (the reference's blocks built unchanged, oracle/_ref) against the golden
vectors the reference itself holds: its two IQ fixtures and the 94-byte frame of tools/debug.txt:5
(committed under tests/golden/ by make_golden.py), plus known answers of the pieces we had to supply
(CRC-32, FFT) and invariance of the mock scheduler to chunking."""
import os
import zlib

import numpy as np
import pytest

import w80211rx

gp, g = w80211rx.gen_ppdu, w80211rx.gen
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(GOLD, name))


def fixture_packets(z):
    post = z["post"]
    return [{"iq": [post[a] for a in range(post.shape[0])], "l_mcs": int(z["l_mcs"]), "l_len": int(z["l_len"]),
             "chan": z["chan"], "cfo": float(z["cfo"]), "seq": 0}]


@pytest.mark.parametrize("name,fmt,mcs", [("fixture_siso.npz", 0, 0), ("fixture_mimo.npz", 1, 8)])
def test_oracle_reproduces_reference_fixture(refgr, name, fmt, mcs):
    """Reference demod/demod2 + decode on the golden post-signal samples -> the frame of tools/debug.txt."""
    z = load(name)
    frame = np.load(os.path.join(GOLD, "frame94.npy")).tobytes()
    nant = z["post"].shape[0]
    streams = refgr.make_sig_stream(fixture_packets(z), nant=nant)
    llr = refgr.run_demod(streams)
    pk = refgr.split_llr_stream(llr)
    assert len(pk) == 1
    assert [pk[0][k] for k in ("format", "mcs", "len", "cr", "ampdu", "trellis", "total")] == list(z["tags"])
    assert np.array_equal(pk[0]["llr"], z["llr"])
    msgs = refgr.run_decode(llr)
    assert len(msgs) == 1
    assert msgs[0] == z["msg"].tobytes()
    assert msgs[0][0] == fmt and msgs[0][-1] == mcs and msgs[0][1] == 94 and msgs[0][2] == 0
    assert msgs[0][3:-1] == frame
    assert zlib.crc32(frame) == 558161692  # the residue decode_impl.cc:374,453 tests for


@pytest.mark.parametrize("chunk_in,chunk_out", [(1, 64), (79, 8192), (8191, 100), (0, 8192)])
def test_mock_scheduler_chunking_invariance(refgr, chunk_in, chunk_out):
    """The reference blocks give the same stream however the scheduler slices their input."""
    z = load("fixture_siso.npz")
    streams = refgr.make_sig_stream(fixture_packets(z) * 2, nant=1)
    llr = refgr.run_demod(streams, chunk_in=chunk_in, chunk_out=chunk_out)
    pk = refgr.split_llr_stream(llr)
    assert len(pk) == 2
    for p in pk:
        assert np.array_equal(p["llr"], z["llr"])
    msgs = refgr.run_decode(llr, chunk_in=max(chunk_in, 0))
    assert msgs == [z["msg"].tobytes()] * 2


def test_fft64_canon_matches_dft():
    """The canonical FFT both sides share is a 64-point DFT to float accuracy (FFTW stand-in)."""
    import ctypes as C
    import subprocess
    import tempfile
    src = r'''
    #include "fft64_canon.h"
    void f(const float* in, float* out) { c8_fft64(in, out); }
    '''
    d = tempfile.mkdtemp()
    open(os.path.join(d, "f.c"), "w").write(src)
    csrc = os.path.join(os.path.dirname(GOLD), "..", "fork-gr-ieee80211_b200", "csrc")
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-I", csrc, os.path.join(d, "f.c"), "-o", os.path.join(d, "f.so")])
    L = C.CDLL(os.path.join(d, "f.so"))
    rng = np.random.default_rng(0)
    x = (rng.standard_normal(64) + 1j * rng.standard_normal(64)).astype(np.complex64)
    y = np.zeros(64, np.complex64)
    L.f(C.c_void_p(x.ctypes.data), C.c_void_p(y.ctypes.data))
    ref = np.fft.fft(x.astype(np.complex128))
    assert np.max(np.abs(y - ref)) < 2e-5 * np.max(np.abs(ref))


def test_viterbi_inverts_encoder(refgr):
    rng = np.random.default_rng(1)
    bits = rng.integers(0, 2, 600, dtype=np.uint8)
    bits[-6:] = 0
    coded = refgr.bcc_encode(bits)
    assert np.array_equal(coded, g.bcc_encode(bits))  # generator's encoder == reference's
    llr = (2.0 * coded - 1.0).astype(np.float32) + 0.3 * rng.standard_normal(coded.size).astype(np.float32)
    assert np.array_equal(refgr.sv_decode(llr, 600), bits)


CASES = [(g.F_L, 0, 1, 40, False), (g.F_L, 5, 1, 333, False), (g.F_HT, 2, 1, 200, False), (g.F_HT, 7, 1, 500, True),
         (g.F_VHT, 4, 1, 260, False), (g.F_VHT, 8, 1, 1000, False), (g.F_HT, 12, 2, 300, False), (g.F_VHT, 7, 2, 512, False)]


@pytest.mark.parametrize("fmt,mcs,nss,length,sgi", CASES)
def test_generator_through_reference_chain(refgr, fmt, mcs, nss, length, sgi):
    """Every generated packet decodes to its own bytes in the reference chain (pins the generator)."""
    b = gp.gen_batch(fmt, mcs, nss, length, 3, snr_db=30, short_gi=sgi, seed=5)
    s = refgr.make_sig_stream(gp.to_packets(b), nant=b["nrx"])
    msgs = refgr.run_decode(refgr.run_demod(s))
    assert [m[3:-1] for m in msgs] == [x.tobytes() for x in b["mpdu"]]


def test_reference_drops(refgr):
    """Corrupted SIG falls back to legacy garbage (FCS fails), short L-SIG duration is dropped by the gate."""
    b1 = gp.gen_batch(g.F_HT, 3, 1, 300, 2, snr_db=30, seed=6, corrupt_sig=True)
    b2 = gp.gen_batch(g.F_VHT, 3, 1, 300, 2, snr_db=30, seed=7, l_len_short=3)
    b3 = gp.gen_batch(g.F_HT, 3, 1, 300, 2, snr_db=30, seed=8, corrupt_fcs=True)
    for b in (b1, b2, b3):
        s = refgr.make_sig_stream(gp.to_packets(b), nant=1)
        assert refgr.run_decode(refgr.run_demod(s)) == []
