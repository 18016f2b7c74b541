"""PPDU assembly on top of gen.py: preamble + SIG fields + data field, optional
2x2 flat channel and AWGN, and the 'post-signal' view that demod / demod2
consume (what the reference's signal_impl.cc:108-192 emits).  Test / bench input
generator only -- not on the hot path."""
import numpy as np

from . import gen as g


def _legacy_preamble(shift):
    stf = np.fft.ifft(g._csd(g._to_fft_bins(g._STF, 26), shift)) * (64.0 / np.sqrt(12))
    ltf = np.fft.ifft(g._csd(g._to_fft_bins(g._L_LTF.astype(complex), 26), shift)) * (64.0 / np.sqrt(52))
    return np.concatenate([np.tile(stf, 3)[:160], ltf[32:], ltf, ltf])


def _nl_ltf(m, iss, n, shift):
    """HT-LTF / VHT-LTF number n of stream iss (19.3.9.4.6, 21.3.8.3.5), 80 samples."""
    pmat = np.array([[1, -1], [1, 1]])
    seq = g._HT_LTF.astype(complex) * pmat[iss][n]
    if m.format == g.F_VHT:
        # pilot tones use the R matrix (row 0 of P) for every stream
        pil = g._PILOT_BINS + 28
        seq[pil] = g._HT_LTF[pil] * pmat[0][n]
    f = g._to_fft_bins(seq, 28)
    return g._ofdm(g._csd(f, shift), 16, 56 * m.nss)


NDP_SIGB_PATTERN = 0b01000010001011100000  # 20 MHz NDP VHT-SIG-B (21.3.8.3.6), bit 0 first; cloud80211phy.cc:1211


def _sigb_symbol(m, shift, ndp=False, mu=False):
    nw = (m.len + 3) // 4
    if ndp:
        b = np.concatenate([g.int_to_bits(NDP_SIGB_PATTERN, 20), np.zeros(6, np.uint8)]).astype(np.uint8)
    elif mu:  # MU: 16-bit length in 4-byte units, 4-bit MCS, tail (Table 21-14)
        b = np.concatenate([g.int_to_bits(nw, 16), g.int_to_bits(m.mcs, 4), np.zeros(6, np.uint8)]).astype(np.uint8)
    else:
        b = np.concatenate([g.int_to_bits(nw, 17), [1, 1, 1], np.zeros(6, np.uint8)]).astype(np.uint8)
    coded = g.bcc_encode(b[None, :])[0]
    inter = g._interleave(coded, g.interleave_perm(52, 1, False, 1))
    f = np.zeros(64, complex)
    f[g._DATA_BINS_NL % 64] = 2.0 * inter - 1.0
    f[g._PILOT_BINS % 64] = np.array([1, 1, 1, -1]) * g.PILOT_POLARITY[3]
    return g._ofdm(g._csd(f, shift), 16, 56 * m.nss), b


def ht_sig_bits(m, short_gi, aggregation=0):
    b = np.zeros(48, np.uint8)
    b[0:7] = g.int_to_bits(m.mcs, 7)
    b[8:24] = g.int_to_bits(m.len, 16)
    b[24] = 1  # smoothing
    b[25] = 1  # not sounding
    b[26] = 1  # reserved
    b[27] = aggregation
    b[31] = 1 if short_gi else 0
    b[34:42] = g.crc8_sig(b[None, :34])[0]
    return b


def vht_siga_bits(m, short_gi):
    b = np.zeros(48, np.uint8)
    b[2] = 1
    b[10:13] = g.int_to_bits(m.nss - 1, 3)
    b[23] = 1
    b[24] = 1 if short_gi else 0
    b[28:32] = g.int_to_bits(m.mcs, 4)
    b[33] = 1
    b[34:42] = g.crc8_sig(b[None, :34])[0]
    return b


def vht_siga_bits_mu(gid, short_gi):
    """VHT-SIG-A of a two-user MU PPDU, one space-time stream per user, BCC for both (Table 21-12)."""
    b = np.zeros(48, np.uint8)
    b[2] = 1
    b[4:10] = g.int_to_bits(gid, 6)
    b[10:13] = g.int_to_bits(1, 3)  # user position 0: one stream
    b[13:16] = g.int_to_bits(1, 3)  # user position 1: one stream
    b[23] = 1
    b[24] = 1 if short_gi else 0
    b[31] = 1  # coding of user 3: reserved 1 when the position is empty
    b[33] = 1
    b[34:42] = g.crc8_sig(b[None, :34])[0]
    return b


def gen_batch(fmt, mcs, nss, length, npkt, snr_db=None, short_gi=False, seed=0, chan=None, nrx=None,
              mpdus=None, corrupt_sig=False, corrupt_fcs=False, l_len_short=0, lead=0, trail=0,
              ndp=False, mu_pos=None, mu_gid=2, psdu=None):
    """Generate npkt packets of one configuration.

    length: PSDU bytes for legacy / HT; MPDU bytes for VHT (wrapped in a one-subframe A-MPDU, so the
            APEP length is length + 4, padded to a multiple of 4).
    chan:   None (stream k straight to antenna k, like the reference's 2x2 fixture) or an
            [nrx, nss] complex matrix (flat channel).
    corrupt_sig: flip one HT-SIG / VHT-SIG-A bit after the CRC was computed (dropped by the SIG check).
    corrupt_fcs: flip one payload bit (packet decodes, FCS fails).
    l_len_short: subtract this many OFDM symbols from the L-SIG duration (dropped by the length gate).
    psdu:   VHT only: a ready-made PSDU [npkt, nbytes] (a multiple of 4 bytes, e.g. from gen.vht_ampdu_multi) sent as it is
            instead of the one-subframe A-MPDU built from `mpdus`.
    ndp:    VHT null data packet (sounding): no data field, the NDP pattern in VHT-SIG-B.
    mu_pos: 0 or 1: a VHT MU PPDU (group id mu_gid) as the user at that position receives it once the other user's
            stream is nulled: legacy part and VHT-SIG-A from chain 0, then VHT-STF, two VHT-LTFs with row mu_pos of the
            P matrix, the MU VHT-SIG-B and one stream of data.  fmt must be VHT, nss 1.
    Returns a dict, see the bottom of this function."""
    rng = np.random.default_rng(seed)
    if mpdus is None:
        mpdus = g.make_mpdus(npkt, length, rng) if psdu is None else np.zeros((npkt, 0), np.uint8)
    if corrupt_fcs:
        mpdus = mpdus.copy()
        mpdus[:, length // 2] ^= 0x10
    if psdu is not None:
        assert fmt == g.F_VHT and psdu.shape[0] == npkt and psdu.shape[1] % 4 == 0
        psdu = np.ascontiguousarray(psdu, np.uint8)
    else:
        psdu = g.vht_ampdu(mpdus) if fmt == g.F_VHT else mpdus
    m = g.Mod(fmt, mcs, nss, psdu.shape[1], short_gi)
    nrx = nrx or m.nss
    streams, bits = g.data_field_bits(m, psdu)
    mu = mu_pos is not None
    if mu:
        assert fmt == g.F_VHT and nss == 1 and not ndp
        m.nltf = 2
    if ndp:
        assert fmt == g.F_VHT

    tx = []
    for iss in range(m.nss):
        sh_l = 4 * iss  # legacy-portion cyclic shift, -200 ns on chain 2
        sh_nl = 8 * (mu_pos if mu else iss)  # HT/VHT-portion cyclic shift, -400 ns on stream 2
        row = mu_pos if mu else iss  # row of the P matrix this stream's LTFs carry
        parts = [np.broadcast_to(_legacy_preamble(sh_l), (npkt, 320))]
        post_len = 0
        if fmt == g.F_L:
            l_mcs, l_len = mcs, length
            data = g._data_symbols(m, streams[iss], iss, sh_l)
        else:
            l_mcs = 0
            nlsym_data = 0 if ndp else -(-(m.nsym * m.nsym_samp) // 80)
            nlsym = 2 + 1 + m.nltf + (1 if fmt == g.F_VHT else 0) + nlsym_data - l_len_short
            l_len = nlsym * 3 - 3
            data = g._data_symbols(m, streams[iss], iss, sh_nl)
        lsig = g._interleave(g.bcc_encode(g._lsig_bits(l_mcs, l_len)[None, :])[0], g.interleave_perm(48, 1, True))
        parts.append(np.broadcast_to(g._sig_symbol(lsig, False, g.PILOT_POLARITY[0], sh_l), (npkt, 80)))
        if fmt != g.F_L:
            sb = ht_sig_bits(m, short_gi) if fmt == g.F_HT else (vht_siga_bits_mu(mu_gid, short_gi) if mu else vht_siga_bits(m, short_gi))
            if corrupt_sig:
                sb = sb.copy()
                sb[9] ^= 1
            c = g.bcc_encode(sb[None, :])[0]
            perm = g.interleave_perm(48, 1, True)
            s1 = g._sig_symbol(g._interleave(c[:48], perm), fmt == g.F_HT, g.PILOT_POLARITY[1], sh_l)
            s2 = g._sig_symbol(g._interleave(c[48:], perm), True, g.PILOT_POLARITY[2], sh_l)
            stf = g._ofdm(g._csd(g._to_fft_bins(g._STF, 26), sh_nl), 16, 12 * m.nss)
            nl = [s1, s2, stf] + [_nl_ltf(m, row, n, sh_nl) for n in range(m.nltf)]
            if fmt == g.F_VHT:
                nl.append(_sigb_symbol(m, sh_nl, ndp=ndp, mu=mu)[0])
            parts.append(np.broadcast_to(np.concatenate(nl), (npkt, sum(len(x) for x in nl))))
        if not ndp:
            parts.append(data)
        tx.append(np.concatenate(parts, axis=1))
    tx = np.stack(tx, axis=0)  # [nss, P, n]

    l_ndbps = g.Mod(g.F_L, l_mcs, 1, l_len).ndbps
    nsamp = -(-(8 * l_len + 22) // l_ndbps) * 80
    total = max(tx.shape[2], 400 + nsamp) + trail
    if chan is None:
        hc = np.eye(nrx, m.nss, dtype=complex)
    else:
        hc = np.asarray(chan, complex).reshape(nrx, m.nss)
    rx = np.zeros((nrx, npkt, lead + total), complex)
    rx[:, :, lead:lead + tx.shape[2]] = np.einsum("rt,tpn->rpn", hc, tx)
    if snr_db is not None:
        sigp = np.mean(np.abs(rx[:, :, lead + 160:lead + 320]) ** 2)
        sd = np.sqrt(sigp / (10.0 ** (snr_db / 10.0)) / 2.0)
        rx = rx + sd * (rng.standard_normal(rx.shape) + 1j * rng.standard_normal(rx.shape))
    rx = rx.astype(np.complex64)

    # what signal_impl hands over: legacy channel from antenna 1 and the samples after L-SIG
    w1 = np.fft.fft(rx[0][:, lead + 160 + 24:lead + 160 + 24 + 64].astype(complex), axis=-1)
    w2 = np.fft.fft(rx[0][:, lead + 160 + 88:lead + 160 + 88 + 64].astype(complex), axis=-1)
    ltf_bins = g._to_fft_bins(g._L_LTF.astype(complex), 26).real
    hl = np.zeros((npkt, 64), np.complex64)
    used = ltf_bins != 0
    hl[:, used] = ((w1[:, used] + w2[:, used]) / (2.0 * ltf_bins[used])).astype(np.complex64)
    post = [np.ascontiguousarray(rx[a][:, lead + 400:lead + 400 + nsamp]) for a in range(nrx)]
    return {
        "mod": m, "format": fmt, "mcs": mcs, "nss": m.nss, "nrx": nrx,
        "mpdu": mpdus, "psdu": psdu, "bits": bits,
        "ppdu": [rx[a] for a in range(nrx)], "sync_idx": lead + 160 + 16,
        "post": post, "l_mcs": l_mcs, "l_len": l_len, "nsamp": nsamp, "chan": hl,
    }


def to_packets(batch, idx=None):
    """gen_batch() output -> list of packet dicts for oracle.refgr.make_sig_stream()."""
    idx = range(batch["post"][0].shape[0]) if idx is None else idx
    return [{"iq": [a[p] for a in batch["post"]], "l_mcs": batch["l_mcs"], "l_len": batch["l_len"],
             "chan": batch["chan"][p], "seq": p} for p in idx]
