"""w80211_viterbi_noalloc (include/w80211rx.h): the decode entry point that allocates nothing, against the existing entry
points on the same LLRs.  No counterpart in the reference.

Named test_zz_* so that pytest collects it last: a failure here must not hide the
parity tests.  Self-contained (own context, own device buffers): passes alone in a fresh process and after the others.
"""
import ctypes as C

import numpy as np
import pytest

import w80211rx

gp, g, host = w80211rx.gen_ppdu, w80211rx.gen, w80211rx.host
pytestmark = pytest.mark.gpu


def descramble_to_bytes(words, trellis, nbytes):
    """decode_impl::descramble + the byte packing of packetAssemble (lib/decode_impl.cc:304-323, 438-449) on the decoded
    trellis bits: the 7-bit scrambler state is read from the first 7 bits, bytes are LSB first from bit 16."""
    bits = np.unpackbits(np.ascontiguousarray(words, "<u4").view(np.uint8), bitorder="little")[:trellis].astype(np.uint8)
    st = [int(b) for b in bits[:7]]  # x1..x7 as they were shifted in
    out = bits.copy()
    out[:7] = 0
    for i in range(7, trellis):
        fb = st[i - 7] ^ st[i - 4]
        st.append(fb)
        out[i] = bits[i] ^ fb
    body = out[16:16 + 8 * nbytes].reshape(-1, 8)
    return (body << np.arange(8, dtype=np.uint8)).sum(axis=1).astype(np.uint8)


def make_case(seed):
    """One packet of every code rate (1/2, 2/3, 3/4, 5/6), legacy / HT / VHT, plus a long one."""
    return [gp.gen_batch(g.F_L, 4, 1, 200, 1, snr_db=25, seed=seed),            # 16-QAM 1/2
            gp.gen_batch(g.F_L, 6, 1, 333, 2, snr_db=28, seed=seed + 1),        # 64-QAM 2/3
            gp.gen_batch(g.F_VHT, 8, 1, 1500, 1, snr_db=30, seed=seed + 2),     # 256-QAM 3/4
            gp.gen_batch(g.F_HT, 7, 1, 4095, 1, snr_db=30, short_gi=True, seed=seed + 3),  # 64-QAM 5/6, trellis 32782 (the maximum)
            gp.gen_batch(g.F_HT, 0, 1, 61, 2, snr_db=12, seed=seed + 4)]        # BPSK 1/2, short and noisy


def run_noalloc(L, torch, dev, desc, llr):
    """-> (list of per-packet word arrays, rc, message) through w80211_viterbi_noalloc, LLRs and bits in device memory."""
    idx = [p for p in range(len(desc)) if desc[p]["status"] == host.ST_OK]
    pk = np.zeros(len(idx), host.NOALLOC_PKT_DTYPE)
    woff = 0
    for k, p in enumerate(idx):
        pk[k] = (int(desc[p]["llr_off"]), woff, int(desc[p]["trellis"]), int(desc[p]["cr"]))
        woff += (int(desc[p]["trellis"]) + 31) // 32
    llr_dev = torch.from_numpy(np.ascontiguousarray(llr)).to(dev)
    bits_dev = torch.full((woff + 8,), 0x5A5A5A5A, dtype=torch.int32, device=dev)  # garbage: every word must be written
    err = C.create_string_buffer(512)
    rc = L.w80211_viterbi_noalloc(dev.index, None, len(pk), pk.ctypes.data, llr_dev.data_ptr(), llr_dev.numel(), bits_dev.data_ptr(),
                                  woff, err, len(err))
    torch.cuda.synchronize(dev)
    w = bits_dev.cpu().numpy().view(np.uint32)
    return idx, [w[int(pk["bits_off"][k]): int(pk["bits_off"][k]) + (int(pk["trellis"][k]) + 31) // 32] for k in range(len(pk))], rc, err.value.decode()


def test_noalloc_entry_equals_the_existing_entry_points_twice_in_a_row():
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    L = host.load()
    rx = host.Receiver(0)  # the existing entry points: w80211_demod_batch for the LLRs, w80211_decode_batch for the bytes
    try:
        for call, seed in enumerate((700, 900)):  # two calls in a row share the static scratch; the second has other packets
            batches = make_case(seed)
            if call == 1:
                batches = batches[::-1]  # other packet order, so the scratch rows hold leftovers of longer packets
            iq, pkts, chan = host.pack_batch(batches)
            dm = rx.demod(iq, pkts, chan)
            ref = rx.decode(dm.desc, dm.llr[: dm.n_llr])
            idx, words, rc, msg = run_noalloc(L, torch, dev, dm.desc, dm.llr[: dm.n_llr])
            assert rc == 0, msg
            assert len(idx) == len(pkts) == 7
            for k, p in enumerate(idx):
                d = dm.desc[p]
                T = int(d["trellis"])
                nbytes = (T - 16) // 8 if int(d["format"]) == host.F_VHT else int(d["len"])
                mine = descramble_to_bytes(words[k], T, nbytes)
                off = int(ref.psdu_off[p])
                theirs = ref.bytes[off: off + nbytes]
                assert np.array_equal(mine, theirs), "call %d packet %d: bytes differ from w80211_decode_batch (bit-exact is the bar)" % (call, p)
            # and the frames the existing path publishes are in those bytes
            assert ref.n_crc_ok >= 5
    finally:
        rx.close()


def test_noalloc_entry_rejects_what_the_scratch_cannot_hold():
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    L = host.load()
    llr = torch.zeros(64 * 17, dtype=torch.float32, device=dev)
    bits = torch.zeros(64, dtype=torch.int32, device=dev)
    err = C.create_string_buffer(512)
    pk = np.zeros(host.NOALLOC_MAX_PKTS + 1, host.NOALLOC_PKT_DTYPE)
    pk["trellis"] = 24
    pk["llr_off"] = np.arange(len(pk)) * 64
    pk["bits_off"] = np.arange(len(pk))
    rc = L.w80211_viterbi_noalloc(0, None, len(pk), pk.ctypes.data, llr.data_ptr(), llr.numel(), bits.data_ptr(), bits.numel(), err, len(err))
    assert rc == -5 and b"W80211_NOALLOC_MAX_PKTS" in err.value
    pk1 = pk[:1].copy()
    pk1["trellis"] = 32783  # beyond W80211_T_MAX
    rc = L.w80211_viterbi_noalloc(0, None, 1, pk1.ctypes.data, llr.data_ptr(), llr.numel(), bits.data_ptr(), bits.numel(), err, len(err))
    assert rc == -5 and b"trellis" in err.value
    pk1["trellis"] = 24
    pk1["llr_off"] = 64 * 17 - 8  # 48 LLRs do not fit
    rc = L.w80211_viterbi_noalloc(0, None, 1, pk1.ctypes.data, llr.data_ptr(), llr.numel(), bits.data_ptr(), bits.numel(), err, len(err))
    assert rc == -1 and b"do not fit" in err.value
    # the accepted maximum works right after the rejections
    ok = pk[: host.NOALLOC_MAX_PKTS].copy()
    rc = L.w80211_viterbi_noalloc(0, None, len(ok), ok.ctypes.data, llr.data_ptr(), llr.numel(), bits.data_ptr(), bits.numel(), err, len(err))
    torch.cuda.synchronize(dev)
    assert rc == 0, err.value
    assert int(bits[: host.NOALLOC_MAX_PKTS].abs().sum()) == 0  # all-zero LLRs: ties keep the even source, the all-zero path
