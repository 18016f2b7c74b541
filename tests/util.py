"""Helpers shared by the parity tests: run the reference chain (oracle/_ref) and compare."""
import numpy as np

import w80211rx

gp = w80211rx.gen_ppdu
host = w80211rx.host

TAG_KEYS = ("format", "mcs", "len", "cr", "ampdu", "trellis", "total")


def ref_chain(refgr, batches, debug=False):
    """-> (list of per-packet dicts from the reference demod's output stream, list of published blobs)."""
    nant = batches[0]["nrx"]
    pk = []
    for b in batches:
        pk += gp.to_packets(b)
    for i, p in enumerate(pk):
        p["seq"] = i
    streams = refgr.make_sig_stream(pk, nant=nant)
    llr = refgr.run_demod(streams)
    pkts = refgr.split_llr_stream(llr)
    msgs = refgr.run_decode(llr, debug=debug)
    return pkts, msgs


def llr_close(a, b, scale):
    """|a-b| <= 1e-4 * max(|a|,|b|,scale): the tolerance of BASELINE.json, SURVEY.md 7.3-5."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.all(np.abs(a - b) <= 1e-4 * np.maximum(np.maximum(np.abs(a), np.abs(b)), scale))


def compare_with_reference(refgr, rx, batches, exact_llr=True, ref=None):
    """Run both sides on the same samples and assert tags, LLRs and published messages agree.
    ref: the reference side's (packets, messages) when the caller has run it already.
    Returns (our Result, reference messages)."""
    iq, pkts, chan = host.pack_batch(batches)
    res = rx.rx(iq, pkts, chan, want_llr=True)
    ref_pkts, ref_msgs = ref if ref is not None else ref_chain(refgr, batches)
    ok_idx = [p for p in range(len(pkts)) if res.desc[p]["status"] == host.ST_OK]
    assert len(ok_idx) == len(ref_pkts), "packets that reach the demod loop: %d vs reference %d" % (len(ok_idx), len(ref_pkts))
    for p, r in zip(ok_idx, ref_pkts):
        d = res.desc[p]
        for k in TAG_KEYS:
            assert int(d[k]) == int(r[k]), "packet %d tag %s: %d vs reference %d" % (p, k, d[k], r[k])
        mine = res.packet_llr(p)
        assert mine.shape == r["llr"].shape
        scale = {1: 1.0, 2: 1.0, 4: 2.0, 6: 4.0, 8: 8.0}[int(d["nbpscs"])]
        assert llr_close(mine, r["llr"], scale), "packet %d LLRs differ by more than 1e-4 relative" % p
        if exact_llr:
            nd = int((mine != r["llr"]).sum())
            assert nd == 0, "packet %d: %d LLRs not bit-identical to the reference" % (p, nd)
        if int(d["format"]) == host.F_VHT:
            assert abs(float(d["sssnr0"]) - float(r["sssnr0"])) <= 1e-3 * max(1.0, abs(float(r["sssnr0"])))
    assert res.messages() == ref_msgs, "published messages differ (%d vs %d)" % (len(res.messages()), len(ref_msgs))
    return res, ref_msgs


def build_stream(batches, rads, gap=300, seed=0, noise=1e-3):
    """Concatenate the raw PPDUs of gen_batch() dicts into one stream per antenna, every packet with its own carrier
    offset (rad per sample), separated by `gap` samples of noise.  -> (raw list, flag positions, rad per flag)"""
    rng = np.random.default_rng(seed)
    nant = batches[0]["nrx"]
    parts = [[] for _ in range(nant)]
    pos, rr = [], []
    off = 0
    k = 0
    for b in batches:
        npkt, n = b["ppdu"][0].shape
        for p in range(npkt):
            rad = float(rads[k % len(rads)])
            k += 1
            lead = (rng.standard_normal((nant, gap)) + 1j * rng.standard_normal((nant, gap))) * noise
            rot = np.exp(-1j * rad * np.arange(n))
            for a in range(nant):
                parts[a].append(lead[a].astype(np.complex64))
                parts[a].append((b["ppdu"][a][p].astype(np.complex128) * rot).astype(np.complex64))
            pos.append(off + gap + b["sync_idx"])
            rr.append(rad)
            off += gap + n
    tail = np.zeros(400, np.complex64)
    raw = [np.concatenate(x + [tail]) for x in parts]
    return raw, np.array(pos, np.int64), np.array(rr, np.float32)


def presiso(x):
    """examples/presiso.grc in numpy: preconj[n] = sum_{k<48} x[n-k-16] conj(x[n-k]), preac[n] = |preconj[n]| / sum_{k<64} |x[n-k]|^2
    (samples before the start of the stream count as zero).  Both sides of a parity test are fed these same arrays."""
    x = np.asarray(x, np.complex128)
    n = len(x)
    d = np.concatenate([np.zeros(16, complex), x[:n - 16]]) * np.conj(x)
    c = np.concatenate([[0], np.cumsum(d)])
    ma = c[1:] - np.concatenate([np.zeros(48, complex), c[1:n - 47]])[:n]
    p = np.concatenate([[0], np.cumsum(np.abs(x) ** 2)])
    pw = p[1:] - np.concatenate([np.zeros(64), p[1:n - 63]])[:n]
    with np.errstate(divide="ignore", invalid="ignore"):
        ac = (np.abs(ma).astype(np.float32) / pw.astype(np.float32)).astype(np.float32)
    return ac, ma.astype(np.complex64)
