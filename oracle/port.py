"""ctypes front end of oracle/libport80211.so (oracle/port/decode_port.c, demod_port.c): the plain-C restatement of the
reference's decode block and of its per-symbol demodulation.  TEST INFRASTRUCTURE ONLY -- see the headers of those files."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libport80211.so")
_lib = None


def available():
    return os.path.exists(_LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_LIB_PATH)
        vp, ci = C.c_void_p, C.c_int
        L.port_viterbi.argtypes = [vp, ci, ci, ci, vp]
        L.port_descramble.argtypes = [vp, vp, ci]
        L.port_trellis_tables.argtypes = [vp, vp]
        L.port_decode_packet.argtypes = [vp, ci, ci, ci, ci, ci, ci, ci, vp, ci, vp, vp]
        L.port_chan_nl.argtypes = [vp, vp]
        L.port_chan_nl.restype = None
        L.port_demod_data.argtypes = [ci, ci, ci, ci, vp, vp, vp]
        L.port_presiso.argtypes = [vp, C.c_long, ci, vp, vp]
        L.port_presiso.restype = None
        _lib = L
    return _lib


def trellis_tables():
    nx, lb = np.zeros(128, np.int32), np.zeros(128, np.int32)
    lib().port_trellis_tables(nx.ctypes.data, lb.ctypes.data)
    return nx.reshape(64, 2), lb.reshape(64, 2)


def viterbi(llr, trellis, cr):
    x = np.ascontiguousarray(llr, np.float32)
    bits = np.zeros(trellis, np.uint8)
    r = lib().port_viterbi(x.ctypes.data, x.size, trellis, cr, bits.ctypes.data)
    if r < 0:
        raise ValueError("port_viterbi: not enough LLRs or bad parameters")
    return bits, r


def decode_packet(p):
    """p: dict with 'llr', 'format', 'mcs', 'len', 'cr', 'ampdu', 'trellis' (the decode block's tags) -> (messages, fcs failures)."""
    x = np.ascontiguousarray(p["llr"], np.float32)
    buf = np.zeros(3 * 4200 + 64, np.uint8)
    nb, nf = C.c_int(0), C.c_int(0)
    n = lib().port_decode_packet(x.ctypes.data, x.size, int(p["format"]), int(p["mcs"]), int(p["len"]), int(p["cr"]),
                                 int(p["ampdu"]), int(p["trellis"]), buf.ctypes.data, buf.size, C.byref(nb), C.byref(nf))
    if n < 0:
        raise ValueError("port_decode_packet failed")
    out, o = [], 0
    for _ in range(n):
        ln = int(buf[o]) | (int(buf[o + 1]) << 8)
        out.append(buf[o + 2:o + 2 + ln].tobytes())
        o += 2 + ln
    return out, nf.value


def run_decode(packets):
    """List of packet dicts (oracle.refgr.split_llr_stream format) -> list of published blobs, in order."""
    msgs = []
    for p in packets:
        msgs += decode_packet(p)[0]
    return msgs


def demod_packet(post, chan_legacy, fmt, nbpscs, nsym, short_gi=False):
    """One packet's post-L-SIG samples on one antenna (what the signal block hands to demod) -> the LLR stream of its
    data symbols (demod_port.c).  fmt 0 legacy / 1 HT / 2 VHT; the caller says what the header stage would have found:
    bits per subcarrier, number of data symbols, short GI.  HT/VHT: the channel comes from the HT/VHT-LTF at sample 240
    (after HT-SIG / VHT-SIG-A and the STF), data starts after it (HT) or after VHT-SIG-B (demod_impl.cc:120,172,199)."""
    x = np.ascontiguousarray(post, np.complex64)
    nsd = 48 if fmt == 0 else 52
    if fmt == 0:
        H = np.ascontiguousarray(chan_legacy, np.complex64)
        base = 0
    else:
        H = np.zeros(64, np.complex64)
        lib().port_chan_nl(x[240:].ctypes.data, H.ctypes.data)
        base = 320 if fmt == 1 else 400
    out = np.zeros(nsym * nsd * nbpscs, np.float32)
    data = x[base:]
    nss = 72 if short_gi else 80
    assert len(data) >= nsym * nss
    r = lib().port_demod_data(fmt, nsym, nss, nbpscs, data.ctypes.data, H.ctypes.data, out.ctypes.data)
    if r < 0:
        raise ValueError("port_demod_data: bad parameters")
    return out


def presiso(x, chunk=0):
    """examples/presiso.grc with GNU Radio's sequential binary32 running sums (oracle/port/presiso_port.c): chunk = 0 carries
    the sums over the whole stream (3.9 / 3.10), chunk > 0 re-makes them every `chunk` outputs (one work() call of 3.7 / 3.8).
    -> (preac float32[n], preconj complex64[n])"""
    x = np.ascontiguousarray(x, np.complex64)
    ac = np.zeros(len(x), np.float32)
    cj = np.zeros(len(x), np.complex64)
    lib().port_presiso(x.ctypes.data, len(x), int(chunk), ac.ctypes.data, cj.ctypes.data)
    return ac, cj
