"""Synthetic 802.11a/g/n/ac 20 MHz PPDU generator (numpy, vectorised over packets).

Input side of the tests and of bench.py: it produces what the reference's
``signal`` / ``signal2`` block hands to ``demod`` / ``demod2`` -- the sample
stream that starts right after L-SIG, the L-SIG rate index and LENGTH, and the
legacy channel estimate (tag ``chan``) -- plus the full PPDU so the reference
chain can be run from the raw samples too.

Written from IEEE 802.11-2016 clauses 17 (OFDM), 19 (HT) and 21 (VHT); the
receiver conventions it must satisfy are the reference's
(/root/reference/lib/demod_impl.cc, demod2_impl.cc, decode_impl.cc), and the
bit pipeline follows the order of /root/reference/lib/encode_impl.cc:130-219
(service, PSDU, tail, pad, scramble with seed 93, BCC, puncture, interleave).
It is NOT part of the hot path and is not timed.
"""
import zlib

import numpy as np

F_L, F_HT, F_VHT = 0, 1, 2
CR_12, CR_23, CR_34, CR_56 = 0, 1, 2, 3

# (nbpscs, cr) per MCS index
_LEGACY = [(1, CR_12), (1, CR_34), (2, CR_12), (2, CR_34), (4, CR_12), (4, CR_34), (6, CR_23), (6, CR_34)]
_HTVHT = [(1, CR_12), (2, CR_12), (2, CR_34), (4, CR_12), (4, CR_34), (6, CR_23), (6, CR_34), (6, CR_56),
          (8, CR_34), (8, CR_56)]
_CR_NUM_DEN = {CR_12: (1, 2), CR_23: (2, 3), CR_34: (3, 4), CR_56: (5, 6)}
_PUNC = {CR_12: [1, 1], CR_23: [1, 1, 1, 0], CR_34: [1, 1, 1, 0, 0, 1], CR_56: [1, 1, 1, 0, 0, 1, 1, 0, 0, 1]}
# L-SIG RATE field R1..R4 (first transmitted first) for 6,9,12,18,24,36,48,54 Mb/s
_LSIG_RATE = [(1, 1, 0, 1), (1, 1, 1, 1), (0, 1, 0, 1), (0, 1, 1, 1), (1, 0, 0, 1), (1, 0, 1, 1), (0, 0, 0, 1), (0, 0, 1, 1)]


class Mod:
    """Per-packet parameters (the reference's c8p_mod, lib/cloud80211phy.h:58-98)."""

    def __init__(self, fmt, mcs, nss, length, short_gi=False):
        self.format, self.mcs, self.nss, self.len = fmt, mcs, nss, length
        self.nsym_samp = 72 if (short_gi and fmt != F_L) else 80
        if fmt == F_L:
            self.nbpscs, self.cr = _LEGACY[mcs]
            self.nsd, self.nss, self.nltf = 48, 1, 0
        else:
            self.nbpscs, self.cr = _HTVHT[mcs % 8] if fmt == F_HT else _HTVHT[mcs]
            self.nsd = 52
            self.nltf = {1: 1, 2: 2}[nss]
        self.ncbpss = self.nbpscs * self.nsd
        self.ncbps = self.ncbpss * self.nss
        num, den = _CR_NUM_DEN[self.cr]
        self.ndbps = self.ncbps * num // den
        nbits = length * 8 + 22
        self.nsym = (nbits + self.ndbps - 1) // self.ndbps
        self.trellis = self.nsym * self.ndbps if fmt == F_VHT else nbits


# ----------------------------------------------------------------- bit level
def scramble_seq(seed, n):
    """x^7+x^4+1, state bit 6 = x7 (reference scramEncoder, cloud80211phy.cc:2594-2605)."""
    out = np.zeros(127, np.uint8)
    st = seed
    for i in range(127):
        fb = ((st >> 6) & 1) ^ ((st >> 3) & 1)
        out[i] = fb
        st = ((st << 1) & 0x7E) | fb
    return np.resize(out, n)


def bcc_encode(bits):
    """K=7 rate-1/2 encoder, generators 133/171 octal; bits [..., n] -> [..., 2n] (A0 B0 A1 B1 ...)."""
    b = np.asarray(bits, np.uint8)
    n = b.shape[-1]
    pad = np.concatenate([np.zeros(b.shape[:-1] + (6,), np.uint8), b], axis=-1)
    d = [pad[..., 6 - k:6 - k + n] for k in range(7)]  # d[k] = input delayed by k
    a = d[0] ^ d[2] ^ d[3] ^ d[5] ^ d[6]  # 133 = 1 011 011
    bb = d[0] ^ d[1] ^ d[2] ^ d[3] ^ d[6]  # 171 = 1 111 001
    out = np.empty(b.shape[:-1] + (2 * n,), np.uint8)
    out[..., 0::2] = a
    out[..., 1::2] = bb
    return out


def puncture(coded, cr):
    pat = np.array(_PUNC[cr], bool)
    n = coded.shape[-1]
    keep = np.resize(pat, n)
    return coded[..., keep]


def interleave_perm(ncbpss, nbpscs, legacy, iss=1):
    """perm such that interleaved[perm[k]] = coded[k] (802.11-2016 17.3.5.7, 19.3.11.8.3)."""
    s = max(nbpscs // 2, 1)
    k = np.arange(ncbpss)
    if legacy:
        ncol, nrow = 16, ncbpss // 16
    else:
        ncol, nrow = 13, 4 * nbpscs
    i = nrow * (k % ncol) + k // ncol
    j = s * (i // s) + (i + ncbpss - (ncol * i) // ncbpss) % s
    if not legacy and iss > 1:
        nrot = 11
        j = (j - ((2 * (iss - 1)) % 3 + 3 * ((iss - 1) // 3)) * nrot * nbpscs) % ncbpss
    return j


def qam_map(bits, nbpscs):
    """Gray mapping of 802.11 (17.3.5.8, 21.3.10.9); bits [..., n*nbpscs] -> complex [..., n], unit mean power."""
    b = np.asarray(bits, np.int64)
    b = b.reshape(b.shape[:-1] + (-1, nbpscs))
    if nbpscs == 1:
        return (2.0 * b[..., 0] - 1.0).astype(np.complex64)
    h = nbpscs // 2

    def axis(bb):  # bb [..., h] MSB-first gray -> level
        g = np.zeros(bb.shape[:-1], np.int64)
        acc = np.zeros(bb.shape[:-1], np.int64)
        for t in range(h):
            acc = acc ^ bb[..., t]  # gray -> binary
            g = (g << 1) | acc
        return 2 * g - (2 ** h - 1)

    re, im = axis(b[..., :h]), axis(b[..., h:])
    scale = {2: 2.0, 4: 10.0, 6: 42.0, 8: 170.0}[nbpscs] ** -0.5
    return ((re + 1j * im) * scale).astype(np.complex64)


def crc8_sig(bits):
    """HT-SIG / VHT-SIG-A CRC (x^8+x^2+x+1, init ones, inverted, c7 first); bits [..., 34] -> [..., 8]."""
    b = np.asarray(bits, np.uint8)
    c = np.full(b.shape[:-1], 0xFF, np.int64)
    for i in range(b.shape[-1]):
        fb = ((c >> 7) & 1) ^ b[..., i]
        c = ((c << 1) & 0xFF) ^ (fb * 0x07)
    c = (~c) & 0xFF
    return np.stack([(c >> (7 - i)) & 1 for i in range(8)], axis=-1).astype(np.uint8)


def bytes_to_bits(by):
    by = np.asarray(by, np.uint8)
    return ((by[..., None] >> np.arange(8)) & 1).reshape(by.shape[:-1] + (-1,)).astype(np.uint8)


def int_to_bits(v, n):
    return np.array([(v >> i) & 1 for i in range(n)], np.uint8)


def make_mpdus(npkt, length, rng):
    """Random MPDUs of `length` bytes whose last 4 bytes are a correct FCS."""
    body = rng.integers(0, 256, size=(npkt, length - 4), dtype=np.uint8)
    out = np.empty((npkt, length), np.uint8)
    out[:, :length - 4] = body
    for p in range(npkt):
        out[p, length - 4:] = np.frombuffer(np.uint32(zlib.crc32(body[p].tobytes())).tobytes(), np.uint8)
    return out


def vht_ampdu(mpdus):
    """One-subframe A-MPDU per packet: 4-byte delimiter (EOF=1, 14-bit length, CRC8, signature 0x4E)
    + MPDU + pad to a multiple of 4 (802.11-2016 9.7.1; the reference's decode walks these at
    decode_impl.cc:336-426 and checks neither CRC nor signature)."""
    npkt, ln = mpdus.shape
    d = np.zeros(16, np.uint8)
    d[0] = 1  # EOF
    d[2] = (ln >> 12) & 1
    d[3] = (ln >> 13) & 1
    for i in range(12):
        d[4 + i] = (ln >> i) & 1
    crc = crc8_sig(d[None, :])[0]
    delim_bits = np.concatenate([d, crc, int_to_bits(0x4E, 8)])
    delim = np.packbits(delim_bits.reshape(4, 8)[:, ::-1], axis=1).reshape(4)
    padn = (-ln) % 4
    out = np.zeros((npkt, 4 + ln + padn), np.uint8)
    out[:, :4] = delim
    out[:, 4:4 + ln] = mpdus
    return out


def vht_ampdu_multi(mpdus, zero_delims=0):
    """A-MPDU of several subframes for ONE packet: `mpdus` is a list of uint8 arrays (an empty array is a zero-length
    delimiter), EOF is set on the last subframe only; `zero_delims` zero-length delimiters (EOF 0) follow every subframe
    but the last.  Returns [1, nbytes].  The reference's walk ORs each delimiter's length into a value that survives
    from the subframe before (+4 after an FCS pass, decode_impl.cc:345-351,414), so consecutive subframes decode cleanly
    only when each length has all the bits of (previous length + 4) set, e.g. lengths growing by 4."""
    parts = []
    for i, m in enumerate(mpdus):
        a = vht_ampdu(np.asarray(m, np.uint8)[None, :])[0].copy()
        last = i == len(mpdus) - 1
        if not last:
            a[0] &= 0xFE
        parts.append(a)
        if not last:
            for _ in range(zero_delims):
                z = vht_ampdu(np.zeros((1, 0), np.uint8))[0].copy()
                z[0] &= 0xFE
                parts.append(z)
    return np.concatenate(parts)[None, :]


# --------------------------------------------------------------- OFDM level
def _seq(s):
    return np.array([int(c) for c in s.replace("-", "n").replace("n1", "-1 ").replace("1", "1 ").replace("0", "0 ").split()])


# L-LTF, subcarriers -26..26 (17.3.3)
_L_LTF = np.array([1, 1, -1, -1, 1, 1, -1, 1, -1, 1, 1, 1, 1, 1, 1, -1, -1, 1, 1, -1, 1, -1, 1, 1, 1, 1, 0,
                   1, -1, -1, 1, 1, -1, 1, -1, 1, -1, -1, -1, -1, -1, 1, 1, -1, -1, 1, -1, 1, -1, 1, 1, 1, 1])
# HT-LTF / VHT-LTF 20 MHz, subcarriers -28..28 (19.3.9.4.6): {1,1,L-LTF,-1,-1}
_HT_LTF = np.concatenate([[1, 1], _L_LTF, [-1, -1]])
# L-STF / HT-STF, subcarriers -26..26, scaled sqrt(1/2)(1+j)
_STF = np.zeros(53, complex)
_STF[[2, 6, 10, 14, 18, 22, 30, 34, 38, 42, 46, 50]] = np.array([1, -1, 1, -1, -1, 1, -1, -1, 1, 1, 1, 1]) * (1 + 1j) / np.sqrt(2)

# pilot polarity sequence p_0..p_126 (17.3.5.10) generated by the scrambler with all-ones seed
PILOT_POLARITY = 1 - 2 * scramble_seq(0x7F, 127).astype(np.int64)

_PILOT_BINS = np.array([-21, -7, 7, 21])
_DATA_BINS_L = np.array([k for k in range(-26, 27) if k != 0 and k not in (-21, -7, 7, 21)])
_DATA_BINS_NL = np.array([k for k in range(-28, 29) if k != 0 and k not in (-21, -7, 7, 21)])


def _to_fft_bins(neg_pos, half):
    """[-half..half] ordered values [..., 2*half+1] -> [..., 64] natural FFT order."""
    out = np.zeros(neg_pos.shape[:-1] + (64,), complex)
    k = np.arange(-half, half + 1)
    out[..., k % 64] = neg_pos
    return out


def _csd(freq, shift_samples):
    """Cyclic shift advance by shift_samples (T_CS = -50 ns * shift): multiply bin k by exp(+j2pi k s/64)."""
    if shift_samples == 0:
        return freq
    k = np.arange(64)
    k = np.where(k < 32, k, k - 64)
    return freq * np.exp(2j * np.pi * k * shift_samples / 64.0)


def _ofdm(freq, cp, ntones):
    """[..., 64] frequency bins -> [..., cp+64] time samples with cyclic prefix."""
    t = np.fft.ifft(freq, axis=-1) * (64.0 / np.sqrt(ntones))
    return np.concatenate([t[..., 64 - cp:], t], axis=-1)


def _sig_symbol(coded48, qbpsk, pol, shift):
    """48 interleaved coded bits -> one 80-sample BPSK/QBPSK symbol on the legacy tone layout."""
    v = (2.0 * coded48.astype(float) - 1.0) * (1j if qbpsk else 1.0)
    f = np.zeros(coded48.shape[:-1] + (64,), complex)
    f[..., _DATA_BINS_L % 64] = v
    f[..., _PILOT_BINS % 64] = np.array([1, 1, 1, -1]) * pol
    return _ofdm(_csd(f, shift), 16, 52)


def _lsig_bits(rate_idx, length):
    b = np.zeros(24, np.uint8)
    b[0:4] = _LSIG_RATE[rate_idx]
    b[5:17] = int_to_bits(length, 12)
    b[17] = b[:17].sum() & 1
    return b


def _interleave(bits, perm):
    out = np.empty_like(bits)
    out[..., perm] = bits
    return out


def data_field_bits(m, psdu):
    """PSDU bytes [P, len] -> interleaved coded bits per stream: list of [P, nsym, ncbpss]."""
    npkt = psdu.shape[0]
    ntot = m.nsym * m.ndbps
    bits = np.zeros((npkt, ntot), np.uint8)
    pb = bytes_to_bits(psdu)
    bits[:, 16:16 + pb.shape[1]] = pb
    scr = scramble_seq(93, ntot)
    if m.format == F_VHT:
        # EOF padding delimiters + pad octets as encode_impl.cc:186-196, tail at the very end
        psdu_len = (ntot - 16 - 6) // 8
        pos = 16 + pb.shape[1]
        eof = np.array([1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0] + list(crc8_sig(np.array([[1] + [0] * 15]))[0]) +
                       list(int_to_bits(0x4E, 8)), np.uint8)
        for _ in range((psdu_len - m.len) // 4):
            bits[:, pos:pos + 32] = eof
            pos += 32
        bits[:, :ntot - 6] ^= scr[:ntot - 6]
    else:
        bits ^= scr
        bits[:, 16 + m.len * 8:16 + m.len * 8 + 6] = 0  # tail
    coded = puncture(bcc_encode(bits), m.cr)  # [P, nsym*ncbps]
    coded = coded.reshape(npkt, m.nsym, m.ncbps)
    legacy = m.format == F_L
    if m.nss == 1:
        return [_interleave(coded, interleave_perm(m.ncbpss, m.nbpscs, legacy, 1))], bits
    s = max(m.nbpscs // 2, 1)
    blk = coded.reshape(npkt, m.nsym, m.ncbpss // s, 2, s)
    streams = []
    for iss in range(2):
        st = blk[:, :, :, iss, :].reshape(npkt, m.nsym, m.ncbpss)
        streams.append(_interleave(st, interleave_perm(m.ncbpss, m.nbpscs, False, iss + 1)))
    return streams, bits


_PSI_1 = np.array([1, 1, 1, -1])
_PSI_2 = [np.array([1, 1, -1, -1]), np.array([1, -1, -1, 1])]


def _data_symbols(m, inter_bits, iss, shift):
    """Interleaved bits [P, nsym, ncbpss] of spatial stream iss -> time samples [P, nsym*nsym_samp]."""
    npkt = inter_bits.shape[0]
    q = qam_map(inter_bits, m.nbpscs)  # [P, nsym, nsd]
    f = np.zeros((npkt, m.nsym, 64), complex)
    n = np.arange(m.nsym)
    if m.format == F_L:
        f[..., _DATA_BINS_L % 64] = q
        pol = PILOT_POLARITY[(n + 1) % 127]
        pil = _PSI_1[None, :] * pol[:, None]
        ntones = 52
    else:
        f[..., _DATA_BINS_NL % 64] = q
        z = 3 if m.format == F_HT else 4
        pol = PILOT_POLARITY[(n + z) % 127]
        if m.format == F_HT and m.nss == 2:
            base = _PSI_2[iss]
        else:
            base = _PSI_1
        idx = (n[:, None] + np.arange(4)[None, :]) % 4
        pil = base[idx] * pol[:, None]
        ntones = 56
    f[..., _PILOT_BINS % 64] = pil[None, :, :]
    if m.nsym_samp == 72 and getattr(m, "sgi_ref", True):
        # The reference takes every FFT window at symbol start + 8 (demod_impl.cc:416), also for
        # 72-sample symbols, while its channel estimate comes from an LTF window that starts 8
        # samples inside a 16-sample CP.  A standard short-GI symbol (8-sample CP) therefore
        # never decodes there.  What it does decode is a long-GI symbol with its last 8 samples
        # cut off; this is the 'short GI' generated by default so that config C3 has FCS-valid
        # packets.  sgi_ref=False gives the standard waveform (both sides then fail the FCS).
        t = _ofdm(_csd(f, shift), 16, ntones * m.nss)[..., :72]
    else:
        t = _ofdm(_csd(f, shift), m.nsym_samp - 64, ntones * m.nss)
    return t.reshape(npkt, m.nsym * m.nsym_samp)
