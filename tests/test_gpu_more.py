"""More GPU parity: the reference's golden fixtures, 2x2, drop gates, ragged / empty batches,
A-MPDUs with several subframes, the split demod / decode entry points, and size-independent
properties at full batch size."""
import ctypes as C
import os
import zlib

import numpy as np
import pytest

import w80211rx
from util import compare_with_reference, ref_chain

gp, g, host = w80211rx.gen_ppdu, w80211rx.gen, w80211rx.host
pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_inputs(name):
    z = np.load(os.path.join(GOLD, name))
    post = z["post"]
    pk = np.zeros(1, host.PKT_IN_DTYPE)
    pk["nsamp"] = post.shape[1]
    pk["l_mcs"] = int(z["l_mcs"])
    pk["l_len"] = int(z["l_len"])
    iq = [np.ascontiguousarray(post[a]) for a in range(post.shape[0])]
    return z, iq, pk, np.ascontiguousarray(z["chan"].reshape(1, 64))


@pytest.mark.parametrize("name", ["fixture_siso.npz", "fixture_mimo.npz"])
def test_reference_fixtures(rx, name):
    """tools/sig80211GenMultiple{Siso_1x1_0,Mimo_2x2_*}.bin -> the 94-byte frame of tools/debug.txt:5."""
    z, iq, pk, chan = golden_inputs(name)
    res = rx.rx(iq, pk, chan, want_llr=True)
    d = res.desc[0]
    assert [int(d[k]) for k in ("format", "mcs", "len", "cr", "ampdu", "trellis", "total")] == list(z["tags"])
    assert np.array_equal(res.packet_llr(0), z["llr"])
    msgs = res.messages()
    assert len(msgs) == 1 and msgs[0] == z["msg"].tobytes()
    assert msgs[0][3:-1] == np.load(os.path.join(GOLD, "frame94.npy")).tobytes()


MIMO = [(g.F_HT, 8 + m, 2, 90 + 61 * m, False) for m in range(8)] + [(g.F_VHT, m, 2, 120 + 75 * m, False) for m in range(9)] \
    + [(g.F_HT, 15, 2, 800, True), (g.F_VHT, 8, 2, 700, True)]


@pytest.mark.parametrize("fmt,mcs,nss,length,sgi", MIMO)
def test_mimo_2x2_all_mcs(refgr, rx, fmt, mcs, nss, length, sgi):
    b = gp.gen_batch(fmt, mcs % 8 + 8 if fmt == g.F_HT else mcs, nss, length, 4, snr_db=32, short_gi=sgi, seed=300 + mcs)
    res, msgs = compare_with_reference(refgr, rx, [b])
    assert len(msgs) == 4


def test_mimo_flat_channel_and_siso_packets_on_two_antennas(refgr, rx):
    """demod2 also takes legacy and 1-SS packets (antenna 1 only) and a mixing channel."""
    h = np.array([[1.0, 0.35 - 0.2j], [0.25 + 0.3j, 0.9 - 0.1j]])
    bs = [gp.gen_batch(g.F_HT, 15, 2, 4095, 2, snr_db=35, seed=11, chan=h),
          gp.gen_batch(g.F_VHT, 5, 2, 600, 2, snr_db=35, seed=12, chan=h),
          gp.gen_batch(g.F_L, 4, 1, 200, 2, snr_db=30, seed=13, nrx=2, chan=h[:, :1]),
          gp.gen_batch(g.F_HT, 5, 1, 300, 2, snr_db=30, seed=14, nrx=2, chan=h[:, :1]),
          gp.gen_batch(g.F_VHT, 6, 1, 300, 2, snr_db=30, seed=15, nrx=2, chan=h[:, :1])]
    res, msgs = compare_with_reference(refgr, rx, bs)
    assert len(msgs) == 10


def test_two_antenna_batch_of_every_symbol_size(refgr, rx):
    """Two-stream packets from BPSK (104 LLRs per symbol) to 256-QAM (832), one-stream packets between them, lengths from
    two symbols to hundreds, in one batch: the demod kernel's staging rows are as long as the batch's largest symbol and a
    block's 16 groups straddle packets of every size."""
    h = np.array([[0.9, 0.3 + 0.25j], [-0.2 + 0.3j, 1.0 - 0.1j]])
    bs = []
    for i, (fmt, mcs, nss, ln) in enumerate([(g.F_HT, 8, 2, 40), (g.F_VHT, 8, 2, 1500), (g.F_HT, 11, 2, 333), (g.F_VHT, 0, 2, 90),
                                             (g.F_HT, 15, 2, 2000), (g.F_VHT, 4, 1, 120), (g.F_L, 0, 1, 60), (g.F_VHT, 7, 2, 64)]):
        bs.append(gp.gen_batch(fmt, mcs, nss, ln, 3, snr_db=36, seed=700 + i, nrx=2, chan=h[:, :nss]))
    res, msgs = compare_with_reference(refgr, rx, bs)
    assert len(msgs) == 24
    # the same packets in the opposite order (the largest symbol no longer first among the two-stream packets)
    res, msgs = compare_with_reference(refgr, rx, bs[::-1])
    assert len(msgs) == 24


def test_drop_gates_and_bad_fcs(refgr, rx):
    """Corrupted SIG, L-SIG duration too short, wrong FCS: the same packets reach the demod loop and the
    same (none) are published; good packets around them are unaffected."""
    bs = [gp.gen_batch(g.F_HT, 3, 1, 300, 2, snr_db=30, seed=20),
          gp.gen_batch(g.F_HT, 3, 1, 300, 2, snr_db=30, seed=21, corrupt_sig=True),
          gp.gen_batch(g.F_VHT, 3, 1, 300, 2, snr_db=30, seed=22, l_len_short=3),
          gp.gen_batch(g.F_HT, 4, 1, 300, 2, snr_db=30, seed=23, l_len_short=2),
          gp.gen_batch(g.F_VHT, 4, 1, 300, 2, snr_db=30, seed=24, corrupt_fcs=True),
          gp.gen_batch(g.F_L, 6, 1, 300, 2, snr_db=30, seed=25, corrupt_fcs=True),
          gp.gen_batch(g.F_L, 3, 1, 77, 2, snr_db=30, seed=26)]
    res, msgs = compare_with_reference(refgr, rx, bs)
    assert len(msgs) == 4
    st = [int(x) for x in res.desc["status"]]
    assert st[4:8] == [host.ST_DROP_GATE] * 4
    assert res.n_mpdu - res.n_crc_ok == 6  # corrupted SIG (decoded as legacy garbage) + wrong FCS


def test_low_snr_fail_set_matches(refgr, rx):
    """At an SNR where some packets fail the FCS, the set of published packets is the reference's."""
    bs = [gp.gen_batch(g.F_HT, 7, 1, 600, 24, snr_db=19.0, seed=31), gp.gen_batch(g.F_VHT, 8, 1, 600, 24, snr_db=24.5, seed=32)]
    res, msgs = compare_with_reference(refgr, rx, bs)
    assert 0 < len(msgs) < 48


def test_ragged_batch_many_lengths(refgr, rx):
    """Packets of every code rate and many lengths in one batch (groups are ragged, tails of every phase)."""
    rng = np.random.default_rng(5)
    bs = []
    for i in range(24):
        fmt = int(rng.integers(0, 3))
        mcs = int(rng.integers(0, 8 if fmt != g.F_VHT else 9))
        ln = int(rng.integers(30, 900))
        bs.append(gp.gen_batch(fmt, mcs, 1, ln, 2, snr_db=30, seed=400 + i))
    res, msgs = compare_with_reference(refgr, rx, bs)
    assert len(msgs) == 48


def test_empty_batch(rx):
    res = rx.rx([np.zeros(16, np.complex64)], np.zeros(0, host.PKT_IN_DTYPE), np.zeros((0, 64), np.complex64))
    assert res.n_mpdu == 0 and res.messages() == []


def test_bad_descriptor_is_flagged_not_crashing(rx):
    b = gp.gen_batch(g.F_L, 3, 1, 100, 2, snr_db=30, seed=40)
    iq, pk, chan = host.pack_batch([b])
    pk["iq_offset"][1] = 10 ** 9  # outside the sample array
    res = rx.rx(iq, pk, chan)
    assert int(res.desc["status"][1]) == host.ST_BAD_INPUT and len(res.messages()) == 1


def test_vht_ampdu_several_subframes(refgr, rx):
    """A-MPDU with three subframes: the delimiter walk of decode_impl.cc:336-426 including its length-accumulation
    quirk (each length has all the bits of the previous length + 4, so all three pass)."""
    rng = np.random.default_rng(9)
    mp = [g.make_mpdus(1, n, rng)[0] for n in (64, 68, 200)]
    b = gp.gen_batch(g.F_VHT, 4, 1, 0, 1, snr_db=30, seed=50, psdu=g.vht_ampdu_multi(mp))
    res, msgs = compare_with_reference(refgr, rx, [b])
    assert res.n_mpdu == 3 and res.n_crc_ok == 3
    assert [m[3:-1] for m in msgs] == [x.tobytes() for x in mp]


def test_vht_ampdu_zero_length_delimiters_between_subframes(refgr, rx):
    """Zero-length delimiters after a subframe: the reference ORs them into the length it carries over, walks the rest
    with that length and publishes only the first subframe; every delimiter it visits is a record here."""
    rng = np.random.default_rng(10)
    mp = [g.make_mpdus(1, n, rng)[0] for n in (64, 68, 72)]
    b = gp.gen_batch(g.F_VHT, 4, 1, 0, 1, snr_db=30, seed=51, psdu=g.vht_ampdu_multi(mp, zero_delims=2))
    res, msgs = compare_with_reference(refgr, rx, [b])
    assert len(msgs) == 1 and res.n_crc_ok == 1 and res.n_mpdu > 1


def test_vht_ampdu_many_subframes_outgrow_the_record_buffers(refgr, rx):
    """36 subframes per PPDU in 8 PPDUs: 288 MPDU records against library buffers that start at 1.5 per packet
    (the post kernel runs once more with the exact count); all 288 are published like the reference does.  Then the same
    batch with out->mpdus = NULL: counted, not returned, no capacity error."""
    rng = np.random.default_rng(11)
    ps = np.concatenate([g.vht_ampdu_multi([g.make_mpdus(1, n, rng)[0] for n in range(28, 28 + 4 * 36, 4)]) for _ in range(8)])
    b = gp.gen_batch(g.F_VHT, 4, 1, 0, 8, snr_db=30, seed=52, psdu=ps)
    res, msgs = compare_with_reference(refgr, rx, [b])
    assert res.n_mpdu == 288 and res.n_crc_ok == 288 and len(msgs) == 288
    iq, pkts, chan = host.pack_batch([b])
    bi = rx.make_in(iq, pkts, chan)
    bo, arrs = rx.make_out_host(len(pkts), 8 * 4096, 1)
    bo.mpdus, bo.mpdu_capacity = None, 0
    rx._check(rx.L.w80211_rx_batch(rx.h, C.byref(bi), C.byref(bo)))
    assert bo.n_mpdu == 288 and bo.n_crc_ok == 288 and bo.psdu_bits_ok == res.psdu_bits_ok
    # a caller buffer that is too small is an error the caller can see, and the context stays usable
    bo2, arrs2 = rx.make_out_host(len(pkts), 8 * 4096, 100)
    assert rx.L.w80211_rx_batch(rx.h, C.byref(bi), C.byref(bo2)) == -5
    res3 = rx.rx(iq, pkts, chan)
    assert res3.n_crc_ok == 288


def test_psdu_of_zero_length_delimiters_only(refgr, rx):
    """A 4 KB PSDU of nothing but zero-length delimiters: about a thousand records per packet, none valid; the reference
    publishes nothing and neither do we (the ADVICE case: records far beyond two per packet)."""
    z = g.vht_ampdu(np.zeros((1, 0), np.uint8))[0].copy()
    z[0] &= 0xFE
    ps = np.tile(z, 1000)[None, :]
    b = gp.gen_batch(g.F_VHT, 8, 1, 0, 1, snr_db=30, seed=53, psdu=ps)
    iq, pkts, chan = host.pack_batch([b])
    bi = rx.make_in(iq, pkts, chan)
    bo, arrs = rx.make_out_host(1, 4096 + 64, 1100)
    rx._check(rx.L.w80211_rx_batch(rx.h, C.byref(bi), C.byref(bo)))
    ref_pkts, ref_msgs = ref_chain(refgr, [b])
    assert ref_msgs == [] and bo.n_crc_ok == 0 and bo.n_mpdu >= 1000  # the symbol padding after the PSDU reads as more of them


def test_split_entry_points_equal_fused(refgr, rx):
    """w80211_demod_batch output feeds w80211_decode_batch (what the two GR blocks do) == fused path;
    decode also accepts the REFERENCE demod's LLR stream."""
    bs = [gp.gen_batch(g.F_L, 7, 1, 1500, 3, snr_db=30, seed=60), gp.gen_batch(g.F_VHT, 7, 1, 900, 3, snr_db=30, seed=61),
          gp.gen_batch(g.F_HT, 5, 1, 123, 3, snr_db=30, seed=62)]
    iq, pk, chan = host.pack_batch(bs)
    fused = rx.rx(iq, pk, chan)
    dm = rx.demod(iq, pk, chan)
    dec = rx.decode(dm.desc, dm.llr[: dm.n_llr])
    assert dec.messages() == fused.messages() and len(fused.messages()) == 9
    ref_pkts, ref_msgs = ref_chain(refgr, bs)
    desc = np.zeros(len(ref_pkts), host.DESC_DTYPE)
    off = 0
    for i, r in enumerate(ref_pkts):
        for k in ("format", "mcs", "len", "cr", "ampdu", "trellis", "total"):
            desc[k][i] = r[k]
        desc["llr_off"][i] = off + 3 * i  # deliberately unaligned offsets
        off += r["total"]
    llr = np.zeros(off + 3 * len(ref_pkts) + 8, np.float32)
    for i, r in enumerate(ref_pkts):
        llr[desc["llr_off"][i]: desc["llr_off"][i] + r["total"]] = r["llr"]
    dec2 = rx.decode(desc, llr)
    assert dec2.messages() == ref_msgs


def test_full_size_batch_properties(rx):
    """BASELINE-size packets in a batch large enough to fill many groups: every packet must come back
    with its own bytes (round trip), the FCS residue must hold, and results must not depend on the slot order."""
    bs = [gp.gen_batch(g.F_VHT, 8, 1, 4088, 96, snr_db=30, seed=70), gp.gen_batch(g.F_HT, 7, 1, 4095, 64, snr_db=30, short_gi=True, seed=71),
          gp.gen_batch(g.F_L, 7, 1, 1500, 200, snr_db=30, seed=72)]
    iq, pk, chan = host.pack_batch(bs)
    res = rx.rx(iq, pk, chan)
    sent = [m.tobytes() for b in bs for m in b["mpdu"]]
    got = res.messages()
    assert [m[3:-1] for m in got] == sent
    assert all(zlib.crc32(m[3:-1]) == 558161692 for m in got)
    perm = np.random.default_rng(1).permutation(len(pk))
    res2 = rx.rx(iq, pk[perm], chan[perm])
    assert [m[3:-1] for m in res2.messages()] == [sent[i] for i in perm]


def test_sharded_by_packet_equals_whole_batch(rx):
    """a batch partitioned by packet (shard.partition, what each rank of a multi-GPU job gets) and
    decoded shard by shard gives, merged in packet order, exactly the messages of the whole batch."""
    shard = w80211rx.shard
    bs = [gp.gen_batch(g.F_L, 6, 1, 700, 5, snr_db=30, seed=61), gp.gen_batch(g.F_HT, 5, 1, 1500, 5, snr_db=30, seed=62),
          gp.gen_batch(g.F_VHT, 7, 1, 400, 5, snr_db=30, seed=63), gp.gen_batch(g.F_HT, 1, 1, 90, 5, snr_db=30, seed=64)]
    iq, pkts, chan = host.pack_batch(bs)
    whole = rx.rx(iq, pkts, chan)
    per_pkt = {}
    for m in whole.mpdus[: whole.n_mpdu]:
        if m["crc_ok"]:
            per_pkt.setdefault(int(m["pkt"]), []).append(whole.bytes[m["off"]: m["off"] + m["len"]].tobytes())
    for world in (2, 3):
        got = {}
        for idx in shard.partition(pkts["nsamp"], world):
            r = rx.rx(iq, pkts[idx], chan[idx])  # same sample arrays, a subset of the packet table
            for m in r.mpdus[: r.n_mpdu]:
                if m["crc_ok"]:
                    got.setdefault(int(idx[m["pkt"]]), []).append(r.bytes[m["off"]: m["off"] + m["len"]].tobytes())
        assert got == per_pkt
    assert len(per_pkt) == 20


@pytest.mark.parametrize("bits", [16, 8])
@pytest.mark.parametrize("nant", [1, 2])
def test_sc16_wire_format(refgr, rx, nant, bits):
    """IQ as int16 pairs + one scale per packet.  The reference is fed the floats a host would
    reconstruct ((float)int16 * scale); the GPU gets the int16 values: same tags, bit-identical LLRs, same messages."""
    if nant == 1:
        bs = [gp.gen_batch(g.F_VHT, 8, 1, 1500, 3, snr_db=30, seed=71), gp.gen_batch(g.F_L, 7, 1, 600, 3, snr_db=30, seed=72),
              gp.gen_batch(g.F_HT, 7, 1, 800, 3, snr_db=30, short_gi=True, seed=73)]
    else:
        bs = [gp.gen_batch(g.F_HT, 15, 2, 900, 3, snr_db=33, seed=74), gp.gen_batch(g.F_VHT, 7, 2, 500, 3, snr_db=33, seed=75)]
    iq, pkts, chan = host.pack_batch(bs)
    q, pk16, back = host.quantize_sc16(iq, pkts, bits=bits)  # bits=8: the sc8 format, same contract
    assert q[0].dtype == (np.int16 if bits == 16 else np.int8) and np.abs(q[0].astype(np.int32)).max() <= (32767 if bits == 16 else 127)
    res = rx.rx(q, pk16, chan, want_llr=True)
    # the same packets as fc32 (the dequantised floats) through the reference blocks
    packets = []
    for p in range(len(pkts)):
        o, n = int(pkts["iq_offset"][p]), int(pkts["nsamp"][p])
        packets.append({"iq": [b[o:o + n] for b in back], "l_mcs": int(pkts["l_mcs"][p]), "l_len": int(pkts["l_len"][p]),
                        "chan": chan[p], "seq": p})
    ref_llr = refgr.run_demod(refgr.make_sig_stream(packets, nant=nant))
    ref_pkts = refgr.split_llr_stream(ref_llr)
    assert len(ref_pkts) == len(pkts)
    for p, r in enumerate(ref_pkts):
        assert np.array_equal(res.packet_llr(p), r["llr"]), "packet %d: LLRs differ" % p
    assert res.messages() == refgr.run_decode(ref_llr)
    assert len(res.messages()) >= (len(pkts) if bits == 16 else len(pkts) - 2)  # 8 bits cost about 0.5 dB at 30 dB SNR
    # and through fc32 input of the same floats: identical to the sc16 path
    res32 = rx.rx(back, pkts, chan, want_llr=True)
    assert np.array_equal(res32.llr[: res32.n_llr], res.llr[: res.n_llr])


def test_garbage_sig_fields_take_the_reference_decisions(refgr, rx):
    """Robustness: the two symbols after L-SIG replaced by noise in 1500 packets (plus random legacy channels for some).
    Format detection then runs on garbage: nearly all fall through to the legacy path and are demodulated as such,
    a few pass a SIG check by chance.  Whatever happens, the same packets must reach the demod loop with the same
    tags, bit-identical LLRs and the same (few or no) published frames."""
    rng = np.random.default_rng(99)
    b = gp.gen_batch(g.F_VHT, 2, 1, 40, 1500, snr_db=25, seed=98)
    post = b["post"][0].copy()
    amp = float(np.sqrt(np.mean(np.abs(post[:, 400:800]) ** 2)))
    post[:, :160] = (amp * (rng.standard_normal((1500, 160)) + 1j * rng.standard_normal((1500, 160))) / np.sqrt(2)).astype(np.complex64)
    b["post"] = [post]
    ch = b["chan"].copy()
    ch[::3] *= np.exp(2j * np.pi * rng.random((len(ch[::3]), 1))).astype(np.complex64)
    b["chan"] = ch
    res, msgs = compare_with_reference(refgr, rx, [b])
    st = np.asarray(res.desc["status"])
    assert (st == host.ST_OK).sum() >= 1400  # the legacy fall-through
    assert len(msgs) <= 5


def test_slots_in_flight_equal_one_batch_at_a_time():
    """w80211_rx_submit / rx_collect with four slots in flight: submit() returns before the header stage has finished and
    the second half of every batch is enqueued by later calls, in submission order; results equal w80211_rx_batch's."""
    rx4 = host.Receiver(0, nslots=4)
    try:
        batches = [gp.gen_batch(g.F_L, 7, 1, 700, 5, snr_db=30, seed=70), gp.gen_batch(g.F_HT, 7, 1, 1200, 4, snr_db=30, short_gi=True, seed=71),
                   gp.gen_batch(g.F_VHT, 8, 1, 2000, 3, snr_db=30, seed=72), gp.gen_batch(g.F_HT, 12, 2, 900, 3, snr_db=35, seed=73),
                   gp.gen_batch(g.F_L, 0, 1, 94, 6, snr_db=20, seed=74), gp.gen_batch(g.F_VHT, 3, 1, 333, 2, snr_db=30, seed=75)]
        packed = [host.pack_batch([b]) for b in batches]
        alone = [rx4.rx(*p, want_llr=True) for p in packed]
        got = [None] * len(packed)
        for rounds in range(2):  # twice: slots are reused
            pend = []
            for i, p in enumerate(packed):
                s = i % 4
                if len(pend) == 4:
                    j = pend.pop(0)
                    got[j] = rx4.collect(j % 4)
                rx4.submit(s, *p, want_llr=True, pinned=rounds == 1)  # page-locked results: nothing in submit() waits for the GPU
                pend.append(i)
            for j in pend:
                got[j] = rx4.collect(j % 4)
            for a, b2, bt in zip(alone, got, batches):
                assert b2.messages() == a.messages() and len(a.messages()) == len(bt["mpdu"])
                assert np.array_equal(b2.desc, a.desc)
                for q in range(len(a.desc)):  # (the gaps between packets in the LLR buffer are not written)
                    assert np.array_equal(b2.packet_llr(q), a.packet_llr(q))
                assert np.array_equal(b2.mpdus[: b2.n_mpdu], a.mpdus[: a.n_mpdu])
    finally:
        rx4.close()


def test_one_survivor_buffer_for_all_slots(monkeypatch):
    """Large batches keep their survivors in one buffer of the context, used by the slots in turn (W80211_SURV_SHARE_BYTES,
    default 1 GiB; 0 here, so that small batches take that path too): batches in flight through rx_submit, and decode-only
    batches in flight through decode_submit (which nothing else orders), some longer than others so that the buffer grows
    while an earlier batch is still using it.  Results equal one batch at a time on a context that does not share."""
    batches = [gp.gen_batch(g.F_L, 7, 1, 300, 6, snr_db=30, seed=270), gp.gen_batch(g.F_VHT, 8, 1, 3000, 40, snr_db=30, seed=271),
               gp.gen_batch(g.F_HT, 7, 1, 1200, 5, snr_db=30, short_gi=True, seed=272), gp.gen_batch(g.F_HT, 15, 2, 4095, 36, snr_db=35, seed=273),
               gp.gen_batch(g.F_L, 0, 1, 94, 3, snr_db=20, seed=274), gp.gen_batch(g.F_VHT, 3, 1, 900, 70, snr_db=30, seed=275)]
    packed = [host.pack_batch([b]) for b in batches]
    plain = host.Receiver(0, nslots=2)
    try:
        alone = [plain.rx(*p, want_llr=True) for p in packed]
    finally:
        plain.close()
    monkeypatch.setenv("W80211_SURV_SHARE_BYTES", "0")
    rx3 = host.Receiver(0, nslots=3)
    try:
        for rounds in range(2):
            got, pend = [None] * len(packed), []
            for i, p in enumerate(packed):
                if len(pend) == 3:
                    j = pend.pop(0)
                    got[j] = rx3.collect(j % 3)
                rx3.submit(i % 3, *p, pinned=True)
                pend.append(i)
            for j in pend:
                got[j] = rx3.collect(j % 3)
            for a, b2, bt in zip(alone, got, batches):
                assert b2.messages() == a.messages() and len(a.messages()) == len(bt["mpdu"])
        # decode-only batches, three in flight, LLRs from the host; page-locked result buffers, so that decode_submit returns
        # while the batch runs (with pageable ones the result copies make it wait) and a batch really is enqueued behind one
        # that still uses the buffer -- without the wait for the previous user's traceback every pair here decodes garbage
        for rounds in range(3):
            got, pend = [None] * len(packed), []
            for i, a in enumerate(alone):
                if len(pend) == 3:
                    j = pend.pop(0)
                    got[j] = rx3.collect(j % 3)
                rx3.decode_submit(i % 3, a.desc, llr=a.llr[: a.n_llr], pinned=True)
                pend.append(i)
            for j in pend:
                got[j] = rx3.collect(j % 3)
            for a, r in zip(alone, got):
                r.desc = a.desc
                assert r.messages() == a.messages()
    finally:
        rx3.close()


def test_submit_collect_errors_leave_the_slot_usable():
    rx2 = host.Receiver(0, nslots=2)
    try:
        b = gp.gen_batch(g.F_L, 5, 1, 400, 3, snr_db=30, seed=80)
        p = host.pack_batch([b])
        with pytest.raises(host.W80211Error):
            rx2.collect(0)  # nothing submitted
        rx2.submit(0, *p)
        with pytest.raises(host.W80211Error):
            rx2.submit(0, *p)  # busy
        ok = rx2.collect(0)
        assert len(ok.messages()) == 3
        # an error of the second half (output buffer too small) is reported by collect(); the slot works again afterwards
        rx2.submit(1, *p, bytes_cap=16)
        with pytest.raises(host.W80211Error):
            rx2.collect(1)
        rx2.submit(1, *p)
        assert rx2.collect(1).messages() == ok.messages()
    finally:
        rx2.close()


def test_llrs_outside_the_stated_domain_never_publish_a_wrong_frame(refgr, rx):
    """DESIGN.md 3 states the Viterbi's input domain: finite LLRs whose path sums stay above the reference's -1e15 sentinel.
    Outside it (NaN / Inf from a zero channel bin, values that overflow binary32) the two decoders need not agree bit for
    bit -- the reference never selects a NaN candidate and keeps the sentinel, fmaxf drops the NaN -- so which of those
    packets still decode may differ; but neither side hangs or faults, neither publishes a frame that was not sent (the
    FCS guards both), hugely scaled but finite LLRs decode like the reference's, and clean packets of the same batch are
    untouched."""
    bs = [gp.gen_batch(g.F_HT, 5, 1, 600, 7, snr_db=30, seed=640)]
    ref_pkts, ref_msgs = ref_chain(refgr, bs)
    assert len(ref_msgs) == 7
    pk = [dict(p) for p in ref_pkts]
    rng = np.random.default_rng(3)
    pk[1]["llr"] = pk[1]["llr"].copy(); pk[1]["llr"][100::52] = np.nan          # one dead subcarrier
    pk[2]["llr"] = pk[2]["llr"].copy(); pk[2]["llr"][rng.integers(0, 5000, 40)] = np.inf
    pk[3]["llr"] = pk[3]["llr"].copy() * np.float32(1e30)                        # finite, far below the sentinel on bad paths
    pk[4]["llr"] = np.full_like(pk[4]["llr"], np.nan)
    pk[5]["llr"] = pk[5]["llr"].copy() * np.float32(3e37)                        # sums overflow to +-Inf, Inf - Inf = NaN
    want = refgr.run_decode(refgr.make_llr_stream(pk))
    desc = np.zeros(len(pk), host.DESC_DTYPE)
    off = 0
    for i, r in enumerate(pk):
        for k in ("format", "mcs", "len", "cr", "ampdu", "trellis", "total"):
            desc[k][i] = r[k]
        desc["llr_off"][i] = off
        off += (r["total"] + 3) & ~3
    llr = np.zeros(off + 8, np.float32)
    for i, r in enumerate(pk):
        llr[desc["llr_off"][i]: desc["llr_off"][i] + r["total"]] = r["llr"]
    got = rx.decode(desc, llr).messages()
    for side in (got, want):
        assert all(m in ref_msgs for m in side) and len(set(side)) == len(side)
        assert ref_msgs[0] in side and ref_msgs[6] in side and ref_msgs[3] in side
        assert ref_msgs[4] not in side
    in_domain = lambda ms: [m for m in ms if m in (ref_msgs[0], ref_msgs[3], ref_msgs[6])]
    assert in_domain(got) == in_domain(want)


def test_demod_and_decode_as_separate_non_blocking_stages_joined_by_llr_leases(monkeypatch):
    """What the two GR blocks do (ABI 7): w80211_demod_submit on one context, the batch's LLRs exported as a lease after
    collect, w80211_decode_submit on ANOTHER context of the same GPU reading that lease -- two batches in flight on each
    side.  Results equal the fused path's; with no byte allowance for older leases (W80211_LEASE_BYTES = 0) a producer keeps
    its W80211_LEASES_KEPT newest, older handles turn invalid and decode refuses them; destroying the producer retires the rest."""
    monkeypatch.setenv("W80211_LEASE_BYTES", "0")
    dm, dc = host.Receiver(0, nslots=2), host.Receiver(0, nslots=2)
    try:
        batches = [gp.gen_batch(g.F_L, 7, 1, 700, 5, snr_db=30, seed=170), gp.gen_batch(g.F_HT, 7, 1, 1200, 4, snr_db=30, short_gi=True, seed=171),
                   gp.gen_batch(g.F_VHT, 8, 1, 2000, 3, snr_db=30, seed=172), gp.gen_batch(g.F_HT, 12, 2, 900, 3, snr_db=35, seed=173)]
        packed = [host.pack_batch([b]) for b in batches]
        fused = [dc.rx(*p) for p in packed]
        leases, descs = [], []
        for i0 in (0, 2):
            for k in (0, 1):
                dm.demod_submit(k, *packed[i0 + k], want_llr=False)
            for k in (0, 1):
                while not dm.poll(k):
                    pass
                r = dm.collect(k)
                descs.append(r.desc.copy())
                leases.append(dm.llr_export(k))
        assert len(set(leases)) == 4 and all(dm.llr_query(h)[0] == 0 for h in leases)
        for i0 in (0, 2):
            for k in (0, 1):
                dc.decode_submit(k, descs[i0 + k], lease=leases[i0 + k])
            for k in (0, 1):
                r = dc.collect(k)
                r.desc = descs[i0 + k]
                assert r.messages() == fused[i0 + k].messages() and len(r.messages()) == len(batches[i0 + k]["mpdu"])
        # host LLRs through the same non-blocking call, decoded from where they are staged (aligned offsets)
        full = dm.demod(*packed[2])
        dc.decode_submit(0, full.desc, llr=full.llr[: full.n_llr])
        r = dc.collect(0)
        r.desc = full.desc
        assert r.messages() == fused[2].messages()
        # more exports than the producer keeps: the oldest handles retire
        for n in range(4):
            dm.demod_submit(0, *packed[0], want_llr=False)
            dm.collect(0)
            leases.append(dm.llr_export(0))
        assert dm.llr_query(leases[0]) is None and dm.llr_query(leases[1]) is None and dm.llr_query(leases[-1]) is not None
        with pytest.raises(host.W80211Error):
            dc.decode_submit(0, descs[0], lease=leases[0])
        dc.decode_submit(0, descs[0], lease=leases[-1])  # the slot is usable after the refusal; same packets as batch 0
        r = dc.collect(0)
        r.desc = descs[0]
        assert r.messages() == fused[0].messages()
        # a reader in flight keeps a lease's memory alive while its producer is destroyed
        dc.decode_submit(1, descs[0], lease=leases[-1])
        dm.close()
        r = dc.collect(1)
        r.desc = descs[0]
        assert r.messages() == fused[0].messages()
        assert dc.llr_query(leases[-1]) is None
    finally:
        dm.close()
        dc.close()


@pytest.mark.parametrize("fmt,mcs,length,sgi", [(0, 7, 1500, False), (1, 7, 4095, True), (2, 8, 4088, False), (2, 3, 77, True)])
def test_llrs_equal_the_plain_c_port(rx, fmt, mcs, length, sgi):
    """The CUDA demod against the plain-C restatement of the reference's per-symbol demodulation (oracle/port/demod_port.c),
    which needs neither the reference tree nor its prebuilt blocks: LLRs of full-size packets, element for element."""
    from oracle import port
    if not port.available():
        pytest.skip("oracle/libport80211.so not built")
    f = [g.F_L, g.F_HT, g.F_VHT][fmt]
    b = gp.gen_batch(f, mcs, 1, length, 4, snr_db=28, short_gi=sgi, seed=950 + mcs)
    iq, pk, chan = host.pack_batch([b])
    res = rx.demod(iq, pk, chan)
    for p in range(len(pk)):
        d = res.desc[p]
        assert int(d["status"]) == host.ST_OK and int(d["format"]) == fmt
        want = port.demod_packet(b["post"][0][p], b["chan"][p], fmt, int(d["nbpscs"]), int(d["nsym"]), short_gi=sgi)
        assert np.array_equal(res.packet_llr(p), want)


def test_one_context_per_device_in_one_process():
    """one ctx per GPU.  Two contexts in one process, on devices 0 and 1 (needs two GPUs: run under
    `gpurun --gpus 2`), each sets its own per-device kernel attributes and decodes its own batch; batches interleaved."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    a, b = host.Receiver(0), host.Receiver(1)
    try:
        bs = [gp.gen_batch(g.F_HT, 15, 2, 1500, 6, snr_db=35, seed=910), gp.gen_batch(g.F_VHT, 8, 1, 3000, 6, snr_db=30, seed=911)]
        packed = [host.pack_batch([x]) for x in bs]
        a.submit(0, *packed[0]); b.submit(0, *packed[1]); a.submit(1, *packed[1]); b.submit(1, *packed[0])
        ra0, rb0, ra1, rb1 = a.collect(0), b.collect(0), a.collect(1), b.collect(1)
        for r, x in ((ra0, bs[0]), (rb1, bs[0]), (rb0, bs[1]), (ra1, bs[1])):
            assert [m[3:-1] for m in r.messages()] == [y.tobytes() for y in x["mpdu"]]
    finally:
        a.close()
        b.close()
