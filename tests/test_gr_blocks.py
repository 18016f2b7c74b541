"""Block-level equivalence: this repo's GNU Radio blocks (gr/lib: demod, demod2, decode -- same factories, io
signatures, tags and messages as the reference's) against the reference's own blocks, both driven by the
same single-threaded mock scheduler over the same streams (oracle/mock_scheduler.h).

The harness library oracle/_ours/libours80211.so = gr/lib/*.cc + the mock GNU Radio surface, linked to the
product libw80211rx.so; oracle/_ref/libref80211.so = the reference's sources, unchanged."""
import os
import sys

import numpy as np
import pytest

import w80211rx
from util import ref_chain, build_stream

gp, g, host = w80211rx.gen_ppdu, w80211rx.gen, w80211rx.host
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OURS_PATH = os.path.join(ROOT, "oracle", "_ours", "libours80211.so")


@pytest.fixture(scope="module")
def ours(refgr):
    assert os.path.exists(OURS_PATH), "oracle/_ours/libours80211.so missing: run __graft_entry__.build() (make -C oracle ours)"
    return refgr.load_harness(OURS_PATH)


@pytest.fixture(autouse=True)
def host_llr_items(monkeypatch):
    """The block-by-block tests compare demod's output items with the reference's, so the LLRs must be in them: the
    hand-off tests below choose their own mode (gr/lib/demod_core.h)."""
    monkeypatch.setenv("W80211_LLR_HANDOFF", "host")


def packets_of(batches):
    pk = []
    for b in batches:
        pk += gp.to_packets(b)
    for i, p in enumerate(pk):
        p["seq"] = i
        p["cfo"], p["snr"], p["rssi"] = 0.01 * i - 0.2, 20.0 + i, -50.0 - i
    return pk


def same_llr_stream(a, b):
    """a, b: Stream objects (possibly of different libraries): same items, same tags on the same items."""
    da, db = a.data(), b.data()
    assert da.shape == db.shape, (da.shape, db.shape)
    assert np.array_equal(da, db), "%d LLRs differ" % int((da != db).sum())
    ta, tb = a.tags(), b.tags()
    assert [o for o, _ in ta] == [o for o, _ in tb]
    for (_, x), (_, y) in zip(ta, tb):
        assert set(x) == set(y), set(x) ^ set(y)
        for k in x:
            if isinstance(x[k], float):
                assert abs(x[k] - y[k]) <= 1e-3 * max(1.0, abs(y[k])), (k, x[k], y[k])
            else:
                assert x[k] == y[k], (k, x[k], y[k])


MIX = [(g.F_L, 0, 1, 94, False), (g.F_L, 7, 1, 1500, False), (g.F_HT, 7, 1, 900, True), (g.F_HT, 2, 1, 200, False),
       (g.F_VHT, 8, 1, 1200, False), (g.F_VHT, 3, 1, 333, True)]


def mixed_batches(nant=1, n=3):
    if nant == 1:
        return [gp.gen_batch(f, m, s, ln, n, snr_db=30, short_gi=sgi, seed=700 + i) for i, (f, m, s, ln, sgi) in enumerate(MIX)] \
            + [gp.gen_batch(g.F_HT, 3, 1, 300, 2, snr_db=30, seed=721, corrupt_sig=True),
               gp.gen_batch(g.F_VHT, 4, 1, 300, 2, snr_db=30, seed=724, corrupt_fcs=True),
               gp.gen_batch(g.F_VHT, 3, 1, 300, 2, snr_db=30, seed=722, l_len_short=3)]
    return [gp.gen_batch(g.F_HT, 15, 2, 1100, n, snr_db=33, seed=800), gp.gen_batch(g.F_VHT, 7, 2, 700, n, snr_db=33, seed=801),
            gp.gen_batch(g.F_HT, 11, 2, 250, n, snr_db=33, short_gi=True, seed=802),
            gp.gen_batch(g.F_L, 5, 1, 200, n, snr_db=30, seed=803, nrx=2),
            gp.gen_batch(g.F_VHT, 2, 1, 150, n, snr_db=30, seed=804, nrx=2)]


@pytest.mark.gpu
@pytest.mark.parametrize("nant", [1, 2])
@pytest.mark.parametrize("chunks", [(0, 8192), (1000, 700), (4099, 8192)])
def test_demod_block_stream_is_the_reference_stream(refgr, ours, nant, chunks):
    pk = packets_of(mixed_batches(nant))
    ref_out = refgr.run_demod(refgr.make_sig_stream(pk, nant=nant))
    our_out = ours.run_demod(ours.make_sig_stream(pk, nant=nant), chunk_in=chunks[0], chunk_out=chunks[1])
    same_llr_stream(our_out, ref_out)


@pytest.mark.gpu
@pytest.mark.parametrize("chunks", [(0, 8192), (777, 8192)])
def test_decode_block_messages_are_the_reference_messages(refgr, ours, chunks):
    """Reference demod -> our decode, our demod -> reference decode, and the all-ours chain."""
    pk = packets_of(mixed_batches(1))
    ref_llr = refgr.run_demod(refgr.make_sig_stream(pk))
    ref_msgs = refgr.run_decode(ref_llr)
    assert len(ref_msgs) >= 18
    ref_pkts = refgr.split_llr_stream(ref_llr)
    assert ours.run_decode(ours.make_llr_stream(ref_pkts), chunk_in=chunks[0]) == ref_msgs
    our_llr = ours.run_demod(ours.make_sig_stream(pk), chunk_in=chunks[0])
    assert ours.run_decode(our_llr, chunk_in=chunks[0]) == ref_msgs
    assert refgr.run_decode(refgr.make_llr_stream(ours.split_llr_stream(our_llr))) == ref_msgs


@pytest.mark.gpu
def test_small_batches_from_the_environment(refgr, ours, monkeypatch):
    """W80211_MAX_BATCH = 2: many GPU calls instead of one; the streams do not change."""
    monkeypatch.setenv("W80211_MAX_BATCH", "2")
    pk = packets_of(mixed_batches(1, n=2))
    ref_llr = refgr.run_demod(refgr.make_sig_stream(pk))
    our_llr = ours.run_demod(ours.make_sig_stream(pk))
    same_llr_stream(our_llr, ref_llr)
    assert ours.run_decode(our_llr) == refgr.run_decode(ref_llr)


def test_chain_runner_equals_block_by_block_on_the_reference(refgr):
    """oracle/mock_scheduler.h mock_run_chain (demod -> decode over a bounded edge, one thread) publishes what the two
    reference blocks publish when each runs over whole streams -- for any edge size and chunking (CPU only)."""
    pk = packets_of([gp.gen_batch(g.F_L, 4, 1, 120, 3, snr_db=30, seed=31), gp.gen_batch(g.F_VHT, 5, 1, 260, 3, snr_db=30, seed=32),
                     gp.gen_batch(g.F_HT, 1, 1, 90, 2, snr_db=30, seed=33, corrupt_fcs=True)])
    want = refgr.run_decode(refgr.run_demod(refgr.make_sig_stream(pk)))
    assert len(want) == 6
    for ring, cin, cout in [(1 << 20, 0, 8192), (16384, 1500, 1024), (16384, 333, 4096)]:
        assert refgr.run_chain(refgr.make_sig_stream(pk), ring_items=ring, chunk_in=cin, chunk_out=cout) == want
        # and with each block on a thread of its own (mock_run_chain_threads)
        assert refgr.run_chain(refgr.make_sig_stream(pk), ring_items=ring, chunk_in=cin, chunk_out=cout, threads=True) == want


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["auto", "device", "both", "host"])
@pytest.mark.parametrize("batch", [(2048, 1), (3, 3), (2, 1)])
@pytest.mark.parametrize("nant", [1, 2])
def test_streaming_chain_in_every_handoff_mode(refgr, ours, monkeypatch, mode, batch, nant):
    """demod -> decode streamed over a bounded edge: batches in flight in both blocks, LLRs handed over in HBM (device: the
    stream items are not even written), on the host, or both; small batches force many leases, several sets per block and
    packets cut by a batch boundary.  The published messages are the reference chain's."""
    monkeypatch.setenv("W80211_LLR_HANDOFF", mode)
    monkeypatch.setenv("W80211_MAX_BATCH", str(batch[0]))
    monkeypatch.setenv("W80211_MIN_BATCH", str(batch[1]))
    pk = packets_of(mixed_batches(nant))
    want = refgr.run_decode(refgr.run_demod(refgr.make_sig_stream(pk, nant=nant)))
    assert len(want) >= 12
    for ring, cin, cout in [(1 << 20, 0, 8192), (20000, 3000, 2048)]:
        got = ours.run_chain(ours.make_sig_stream(pk, nant=nant), ring_items=ring, chunk_in=cin, chunk_out=cout)
        assert got == want, (ring, len(got), len(want))
        # each block on a thread of its own, as under GNU Radio's scheduler: two contexts driven from two threads, the lease
        # table shared between them
        got = ours.run_chain(ours.make_sig_stream(pk, nant=nant), ring_items=ring, chunk_in=cin, chunk_out=cout, threads=True)
        assert got == want, ("threads", ring, len(got), len(want))


@pytest.mark.gpu
def test_device_handoff_writes_no_llr_items(refgr, ours, monkeypatch):
    """W80211_LLR_HANDOFF=device: demod's stream has the reference's tags plus the lease tags, `total` items per packet,
    and the items are left unwritten; `both` carries the reference's LLRs as well.  A decode block that finds the lease
    retired falls back to the items when they exist and drops the packet (and says so) when they do not."""
    pk = packets_of(mixed_batches(1, n=2))
    ref_llr = refgr.run_demod(refgr.make_sig_stream(pk))
    monkeypatch.setenv("W80211_LLR_HANDOFF", "both")
    both = ours.run_demod(ours.make_sig_stream(pk))
    assert np.array_equal(both.data(), ref_llr.data())
    monkeypatch.setenv("W80211_LLR_HANDOFF", "device")
    dev = ours.run_demod(ours.make_sig_stream(pk))
    assert dev.data().shape == ref_llr.data().shape and not dev.data().any()
    for s in (both, dev):
        for (o, x), (o2, y) in zip(s.tags(), ref_llr.tags()):
            assert o == o2 and set(x) - set(y) <= {"w80211_llr", "w80211_off", "w80211_nohost"} and set(y) <= set(x)
            assert x["w80211_llr"] > 0 and x["w80211_off"] % 4 == 0 and ("w80211_nohost" in x) == (s is dev)
    # the demod blocks are gone and with them their leases: `both` still decodes from the items, `device` cannot
    want = refgr.run_decode(ref_llr)
    assert ours.run_decode(both) == want
    assert ours.run_decode(dev) == []


def _capture_stdout(fn):
    import tempfile
    sys.stdout.flush()
    with tempfile.TemporaryFile() as tf:
        old = os.dup(1)
        os.dup2(tf.fileno(), 1)
        try:
            r = fn()
        finally:
            os.dup2(old, 1)
            os.close(old)
        tf.seek(0)
        return r, tf.read().decode()


@pytest.mark.gpu
def test_decode_debug_lines(refgr, ours):
    """ifdebug = true: the same 'crc32 correct / wrong' lines with the same counters (decode_impl.cc:376-509)."""
    pk = packets_of(mixed_batches(1, n=2))
    ref_llr = refgr.run_demod(refgr.make_sig_stream(pk))
    ref_pkts = refgr.split_llr_stream(ref_llr)
    m1, t1 = _capture_stdout(lambda: refgr.run_decode(refgr.make_llr_stream(ref_pkts), debug=True))
    m2, t2 = _capture_stdout(lambda: ours.run_decode(ours.make_llr_stream(ref_pkts), debug=True))
    assert m1 == m2
    l1 = [x for x in t1.splitlines() if x.startswith("ieee80211 decode")]
    l2 = [x for x in t2.splitlines() if x.startswith("ieee80211 decode")]
    assert l1 and l1 == l2


def test_blocks_without_a_gpu_drop_packets_and_say_so(refgr, capfd):
    """No CUDA device: the blocks log the failure, consume their input and emit nothing -- no CPU path, no exception
    out of general_work (reference blocks never throw either)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    if not os.path.exists(OURS_PATH):
        pytest.skip("harness not built")
    ours = refgr.load_harness(OURS_PATH)
    pk = packets_of([gp.gen_batch(g.F_L, 3, 1, 100, 2, snr_db=30, seed=1)])
    s = ours.make_sig_stream(pk)
    out = ours.run_demod(s)
    assert out.data().size == 0 and out.tags() == []
    assert "no usable CUDA device" in capfd.readouterr().err
    # the detect block passes its samples unflagged: one zero byte per sample, no tags
    x = (np.random.default_rng(0).standard_normal(5000) + 0j).astype(np.complex64)
    flags = ours.run_detect(x, chunk_in=1500)
    assert flags.data().size == 5000 and not flags.data().any() and flags.tags() == []
    assert "no usable CUDA device" in capfd.readouterr().err


# ------------------------------------------------------------------ signal / signal2
def same_sig_streams(ours_out, ref_out):
    """Packets at the same items with the same tags; samples and channel to 1e-4 of the packet's peak (the rotation
    factors come from two different cosf/sinf, see tests/test_gpu_signal.py); pad items are not compared (the
    reference leaves them unwritten)."""
    to, tr = ours_out[0].tags(), ref_out[0].tags()
    assert [o for o, _ in to] == [o for o, _ in tr]
    for a in range(1, len(ref_out)):
        assert ours_out[a].tags() == [] and ref_out[a].tags() == []
    do, dr = [s.data() for s in ours_out], [s.data() for s in ref_out]
    for a in range(len(dr)):
        assert do[a].shape == dr[a].shape
    for (off, x), (_, y) in zip(to, tr):
        assert set(x) == set(y)
        for k in ("seq", "mcs", "len", "nsamp"):
            assert x[k] == y[k], (k, x[k], y[k])
        for k in ("cfo", "snr", "rssi"):
            assert np.float32(x[k]) == np.float32(y[k]), (k, x[k], y[k])
        assert np.all(np.abs(x["chan"] - y["chan"]) <= 1e-4 * np.abs(y["chan"]).max())
        n = y["nsamp"]
        for a in range(len(dr)):
            ref = dr[a][off:off + n]
            assert np.all(np.abs(do[a][off:off + n] - ref) <= 1e-4 * np.abs(ref).max()), "packet at %d antenna %d" % (off, a)
    return len(tr)


@pytest.mark.gpu
@pytest.mark.parametrize("nant", [1, 2])
@pytest.mark.parametrize("chunks", [(0, 8192), (1000, 700), (4099, 8192)])
def test_signal_block_stream_matches_the_reference(refgr, ours, nant, chunks):
    if nant == 1:
        bs = [gp.gen_batch(g.F_L, 5, 1, 300, 2, snr_db=30, seed=900), gp.gen_batch(g.F_HT, 6, 1, 800, 2, snr_db=30, seed=901),
              gp.gen_batch(g.F_VHT, 7, 1, 1000, 2, snr_db=30, seed=902)]
    else:
        bs = [gp.gen_batch(g.F_HT, 11, 2, 600, 2, snr_db=35, seed=903), gp.gen_batch(g.F_VHT, 4, 2, 900, 2, snr_db=35, seed=904)]
    raw, pos, rad = build_stream(bs, [0.0071, -0.0123, 0.0004], gap=500, noise=0.02)
    # a false flag on noise before every packet and one inside every packet
    allpos = np.sort(np.concatenate([pos, pos - 400, pos + 1000]))
    rads = np.repeat(rad, 3)
    ref_out = refgr.run_signal(raw, [int(x) for x in allpos], rad=rads, snr=21.5, rssi=-40.0)
    our_out = ours.run_signal(raw, [int(x) for x in allpos], rad=rads, snr=21.5, rssi=-40.0, chunk_in=chunks[0], chunk_out=chunks[1])
    assert same_sig_streams(our_out, ref_out) == len(pos)
    # the rest of the chain on each side's own stream: same published messages
    ref_msgs = refgr.run_decode(refgr.run_demod(ref_out))
    our_msgs = ours.run_decode(ours.run_demod(our_out))
    assert our_msgs == ref_msgs and len(ref_msgs) >= len(pos) - 1


# ------------------------------------------------------------------ detect
@pytest.mark.gpu
@pytest.mark.parametrize("chunks", [(0, 8192, 1 << 20), (1000, 700, 1 << 20), (4099, 8192, 2048), (0, 8192, 700)])
def test_detect_block_flags_are_the_reference_chain_s(refgr, ours, monkeypatch, chunks):
    """This repo's `detect` block (one block for presiso + trigger + sync) against the reference's trigger -> sync blocks fed
    GNU Radio's presiso streams (oracle/port/presiso_port.c, running sums carried as in 3.10): a flag at the same items, with
    rad / snr / rssi to the tolerance of the presiso contract, for scheduler windows and GPU chunks of several sizes (chunks
    that cut runs, countdowns and sync windows); then the rest of the chain on each side's flags publishes the same frames."""
    from oracle import port
    monkeypatch.setenv("W80211_DETECT_CHUNK", str(chunks[2]))
    bs = [gp.gen_batch(g.F_HT, 5, 1, 300, 3, snr_db=25, seed=950), gp.gen_batch(g.F_L, 3, 1, 100, 3, snr_db=25, seed=951),
          gp.gen_batch(g.F_VHT, 4, 1, 500, 3, snr_db=25, seed=952)]
    raw, pos, rad = build_stream(bs, [0.01, -0.02, 0.004], gap=600, noise=0.03, seed=95)
    x = raw[0]
    ac, cj = port.presiso(x, 0)
    ref = refgr.run_sync(refgr.run_trigger(ac), cj, x)
    mine = ours.run_detect(x, chunk_in=chunks[0], chunk_out=chunks[1])
    rt, mt = ref.tags(), mine.tags()
    assert [o for o, _ in mt] == [o for o, _ in rt] and len(rt) == len(pos)
    md, rd = mine.data(), ref.data()
    assert len(md) <= len(x) and np.array_equal(np.flatnonzero(md), np.flatnonzero(rd[: len(md)]))
    for (_, a), (_, b) in zip(mt, rt):
        assert set(a) == set(b) == {"rad", "snr", "rssi"}
        assert abs(a["rad"] - b["rad"]) <= 1e-4 and abs(a["snr"] - b["snr"]) <= 1e-2 * max(1.0, abs(b["snr"]))
        assert abs(a["rssi"] - b["rssi"]) <= 1e-4 * abs(b["rssi"])
    ref_msgs = refgr.run_decode(refgr.run_demod(refgr.run_signal(raw, [o for o, _ in rt], rad=[t["rad"] for _, t in rt])))
    sig = ours.Block("signal")
    out = ours.Stream(itemsize=8)
    sig.run([mine, ours.Stream(x[: len(md)])], [out])
    assert ours.run_decode(ours.run_demod([out])) == ref_msgs and len(ref_msgs) == len(pos)
