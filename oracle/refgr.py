"""ctypes front end for oracle/_ref/libref80211.so -- the reference's own block
sources (signal, signal2, demod, demod2, decode + cloud80211phy), compiled
unchanged against oracle/mockgr and driven by oracle/mock_scheduler.h.

TEST INFRASTRUCTURE ONLY.  Importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs; never from the product path.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_ref", "libref80211.so")
_lib = None


def available():
    return os.path.exists(_LIB_PATH)


def load_harness(path):
    """A second copy of this module bound to another library with the same stream / block entry points:
    oracle/_ours/libours80211.so, this repo's GNU Radio blocks under the same mock scheduler."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("refgr_harness_" + os.path.basename(path).replace(".", "_"), __file__)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod._LIB_PATH = path
    return mod


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_LIB_PATH)
        vp, ci, cl, cf, cd = C.c_void_p, C.c_int, C.c_long, C.c_float, C.c_double
        L.ref_stream_new.restype = vp
        L.ref_stream_new.argtypes = [ci, cl, vp]
        L.ref_stream_free.argtypes = [vp]
        L.ref_stream_nitems.restype = cl
        L.ref_stream_nitems.argtypes = [vp]
        L.ref_stream_itemsize.argtypes = [vp]
        L.ref_stream_data.restype = vp
        L.ref_stream_data.argtypes = [vp]
        L.ref_stream_add_tag_long.argtypes = [vp, cl, C.c_char_p, cl]
        L.ref_stream_add_tag_float.argtypes = [vp, cl, C.c_char_p, cf]
        L.ref_stream_add_tag_c32.argtypes = [vp, cl, C.c_char_p, vp, ci]
        L.ref_stream_sort_tags.argtypes = [vp]
        L.ref_chain_run.restype = cl
        L.ref_chain_run.argtypes = [vp, vp, C.POINTER(vp), ci, cl, cl, cl, ci]
        L.ref_stream_ntags.argtypes = [vp]
        L.ref_stream_tag_offset.restype = cl
        L.ref_stream_tag_offset.argtypes = [vp, ci]
        L.ref_stream_tag_key.restype = C.c_char_p
        L.ref_stream_tag_key.argtypes = [vp, ci]
        L.ref_stream_tag_kind.argtypes = [vp, ci]
        L.ref_stream_tag_long.restype = cl
        L.ref_stream_tag_long.argtypes = [vp, ci]
        L.ref_stream_tag_double.restype = cd
        L.ref_stream_tag_double.argtypes = [vp, ci]
        L.ref_stream_tag_c32.argtypes = [vp, ci, vp, ci]
        L.ref_block_new.restype = vp
        L.ref_block_new.argtypes = [C.c_char_p, ci, ci]
        L.ref_block_free.argtypes = [vp]
        L.ref_block_run.restype = cl
        L.ref_block_run.argtypes = [vp, C.POINTER(vp), ci, C.POINTER(vp), ci, cl, cl]
        L.ref_block_nmsgs.argtypes = [vp]
        L.ref_block_msg_len.argtypes = [vp, ci]
        L.ref_block_msg_data.restype = vp
        L.ref_block_msg_data.argtypes = [vp, ci]
        L.ref_block_clear_msgs.argtypes = [vp]
        if not hasattr(L, "ref_parse_legacy"):  # the harness build of this repo's own blocks (oracle/_ours) has no PHY entry points
            _lib = L
            return _lib
        L.ref_parse_legacy.argtypes = [ci, ci, vp]
        L.ref_parse_ht.argtypes = [vp, vp]
        L.ref_parse_vht.argtypes = [vp, vp, vp]
        L.ref_qam_to_llr.argtypes = [ci, ci, vp, vp]
        L.ref_deint.argtypes = [ci, ci, vp, vp]
        L.ref_depas2.argtypes = [ci, ci, vp, vp, vp]
        L.ref_sv_decode.argtypes = [vp, vp, ci]
        L.ref_bcc_encode.argtypes = [vp, vp, ci]
        L.ref_sv_tables.argtypes = [vp, vp]
        L.ref_pilot_p.argtypes = [vp]
        L.ref_table.argtypes = [C.c_char_p, vp, ci]
        L.ref_intel_vhtb20.argtypes = [vp, vp]
        L.ref_check_crc8.argtypes = [vp]
        L.ref_nlsig_demod.argtypes = [vp, vp, vp, vp, vp]
        _lib = L
    return _lib


_DTYPES = {1: np.uint8, 4: np.float32, 8: np.complex64}


class Stream:
    """A finite GNU Radio stream: items + tags {offset: {key: value}}."""

    def __init__(self, data=None, itemsize=None):
        L = lib()
        if data is None:
            self.h = L.ref_stream_new(int(itemsize), 0, None)
        else:
            a = np.ascontiguousarray(data)
            self.h = L.ref_stream_new(a.dtype.itemsize, a.size, a.ctypes.data)

    def __del__(self):
        try:
            lib().ref_stream_free(self.h)
        except Exception:
            pass

    def add_tag(self, offset, key, value):
        L = lib()
        k = key.encode()
        if isinstance(value, (int, np.integer)):
            L.ref_stream_add_tag_long(self.h, int(offset), k, int(value))
        elif isinstance(value, (float, np.floating)):
            L.ref_stream_add_tag_float(self.h, int(offset), k, float(value))
        else:
            v = np.ascontiguousarray(value, dtype=np.complex64)
            L.ref_stream_add_tag_c32(self.h, int(offset), k, v.ctypes.data, v.size)

    def data(self):
        L = lib()
        n = L.ref_stream_nitems(self.h)
        isz = L.ref_stream_itemsize(self.h)
        dt = _DTYPES[isz]
        if n == 0:
            return np.zeros(0, dt)
        buf = (C.c_char * (n * isz)).from_address(L.ref_stream_data(self.h))
        return np.frombuffer(buf, dtype=dt).copy()

    def tags(self):
        """-> list of (offset, dict) sorted by offset, one dict per tagged item."""
        L = lib()
        out = {}
        for i in range(L.ref_stream_ntags(self.h)):
            off = L.ref_stream_tag_offset(self.h, i)
            key = L.ref_stream_tag_key(self.h, i).decode()
            kind = L.ref_stream_tag_kind(self.h, i)
            if kind == 0:
                val = int(L.ref_stream_tag_long(self.h, i))
            elif kind == 1:
                val = float(np.float32(L.ref_stream_tag_double(self.h, i)))
            elif kind == 2:
                n = L.ref_stream_tag_c32(self.h, i, None, 0)
                v = np.zeros(n, np.complex64)
                L.ref_stream_tag_c32(self.h, i, v.ctypes.data, n)
                val = v
            else:
                val = None
            out.setdefault(off, {})[key] = val
        return sorted(out.items())


class Block:
    """One reference block instance: kind in signal, signal2, demod, demod2, decode."""

    def __init__(self, kind, a0=0, a1=0):
        self.kind = kind
        self.h = lib().ref_block_new(kind.encode(), int(a0), int(a1))
        if not self.h:
            raise ValueError(kind)

    def __del__(self):
        try:
            lib().ref_block_free(self.h)
        except Exception:
            pass

    def run(self, ins, outs=None, chunk_in=0, chunk_out=8192):
        L = lib()
        if outs is None:
            outs = []
        elif isinstance(outs, Stream):
            outs = [outs]
        arr = (C.c_void_p * len(ins))(*[s.h for s in ins])
        oarr = (C.c_void_p * max(len(outs), 1))(*[s.h for s in outs])
        r = L.ref_block_run(self.h, arr, len(ins), oarr, len(outs), chunk_in, chunk_out)
        if r < 0:
            raise RuntimeError("reference block raised")
        return r

    def messages(self, clear=True):
        L = lib()
        msgs = []
        for i in range(L.ref_block_nmsgs(self.h)):
            n = L.ref_block_msg_len(self.h, i)
            buf = (C.c_char * n).from_address(L.ref_block_msg_data(self.h, i))
            msgs.append(bytes(buf))
        if clear:
            L.ref_block_clear_msgs(self.h)
        return msgs


# ------------------------------------------------------------------ chains
def run_signal(iq_list, sync_idx, rad=0.0, snr=0.0, rssi=0.0, chunk_in=0, chunk_out=8192):
    """Raw PPDU samples + sync positions -> stream(s) in demod's input format.
    iq_list: one (SISO) or two (2x2) complex64 arrays. sync_idx: indices of
    'L-LTF start + 16' (reference lib/sync_impl.cc:123)."""
    n = len(iq_list[0])
    sync = np.zeros(n, np.uint8)
    s_sync = None
    sync_idx = list(np.atleast_1d(sync_idx))
    rads = np.broadcast_to(np.asarray(rad, np.float32), (len(sync_idx),))
    for i in sync_idx:
        sync[i] = 1
    s_sync = Stream(sync)
    for i, r in zip(sync_idx, rads):
        s_sync.add_tag(i, "rad", float(r))
        s_sync.add_tag(i, "snr", float(snr))
        s_sync.add_tag(i, "rssi", float(rssi))
    ins = [s_sync] + [Stream(np.asarray(x, np.complex64)) for x in iq_list]
    blk = Block("signal" if len(iq_list) == 1 else "signal2")
    outs = [Stream(itemsize=8) for _ in iq_list]
    blk.run(ins, outs, chunk_in, chunk_out)
    return outs


def run_trigger(preac, chunk_in=0, chunk_out=8192):
    """presiso's autocorrelation stream -> the trigger block's flag stream (uint8: 0x01 trigger, 0x02 new maximum)."""
    blk = Block("trigger")
    out = Stream(itemsize=1)
    blk.run([Stream(np.asarray(preac, np.float32))], out, chunk_in, chunk_out)
    return out


def run_sync(trigger_stream, preconj, sig, chunk_in=0, chunk_out=8192):
    """trigger flags + presiso's conjugate stream + samples -> the sync block's flag stream with tags rad / snr / rssi."""
    blk = Block("sync")
    out = Stream(itemsize=1)
    blk.run([trigger_stream, Stream(np.asarray(preconj, np.complex64)), Stream(np.asarray(sig, np.complex64))], out, chunk_in, chunk_out)
    return out


def run_detect(sig, chunk_in=0, chunk_out=8192):
    """received samples -> the sync flag stream with tags rad / snr / rssi, through this repo's `detect` block (the harness
    build of gr/lib only: the reference has presiso + trigger + sync in its place)."""
    blk = Block("detect")
    out = Stream(itemsize=1)
    blk.run([Stream(np.asarray(sig, np.complex64))], out, chunk_in, chunk_out)
    return out


def run_demod(sig_streams, mupos=0, mugid=2, chunk_in=0, chunk_out=8192):
    """signal-format stream(s) -> LLR stream with demod's tags."""
    blk = Block("demod", mupos, mugid) if len(sig_streams) == 1 else Block("demod2")
    out = Stream(itemsize=4)
    blk.run(sig_streams, out, chunk_in, chunk_out)
    return out


def run_decode(llr_stream, debug=False, chunk_in=0, chunk_out=8192):
    """LLR stream with tags -> list of published blobs (bytes)."""
    blk = Block("decode", 1 if debug else 0)
    blk.run([llr_stream], None, chunk_in, chunk_out)
    return blk.messages()


def run_chain(sig_streams, mupos=0, mugid=2, debug=False, ring_items=1 << 20, chunk_in=0, chunk_out=8192, blocks=None, timing=None,
              threads=False):
    """demod / demod2 -> decode joined by a bounded edge and run in turn by one thread (mock_run_chain), the way a
    flowgraph streams: signal-format stream(s) -> list of published blobs.  threads: each block on a thread of its own
    (mock_run_chain_threads), like under GNU Radio's thread-per-block scheduler.  blocks: an (upstream, downstream) pair to
    reuse (both must stand at a packet boundary); timing: a dict that receives the seconds spent inside the two blocks'
    run ("chain_s") and the general_work calls made."""
    import time
    L = lib()
    if blocks is None:
        blocks = chain_blocks(len(sig_streams), mupos, mugid, debug)
    up, down = blocks
    arr = (C.c_void_p * len(sig_streams))(*[s.h for s in sig_streams])
    t0 = time.perf_counter()
    r = L.ref_chain_run(up.h, down.h, arr, len(sig_streams), ring_items, chunk_in, chunk_out, 1 if threads else 0)
    if timing is not None:
        timing["chain_s"] = time.perf_counter() - t0
        timing["calls"] = int(r)
    if r < 0:
        raise RuntimeError("block raised")
    return down.messages()


def chain_blocks(nant=1, mupos=0, mugid=2, debug=False):
    return (Block("demod", mupos, mugid) if nant == 1 else Block("demod2")), Block("decode", 1 if debug else 0)


def make_sig_stream(packets, nant=1):
    """Build demod-input stream(s) from packet dicts:
    {'iq': [ant0, ant1?] complex64 arrays of nsamp samples, 'l_mcs', 'l_len', 'chan': c64[64],
     'cfo','snr','rssi','seq'}.  320 zero pad items follow each packet, as
    signal_impl.cc:194-201 reserves (unwritten there)."""
    bufs = [[] for _ in range(nant)]
    tags = []
    off = 0
    for k, p in enumerate(packets):
        nsamp = len(p["iq"][0])
        tags.append((off, p, nsamp, k))
        for a in range(nant):
            bufs[a].append(np.asarray(p["iq"][a], np.complex64))
            bufs[a].append(np.zeros(320, np.complex64))
        off += nsamp + 320
    streams = [Stream(np.concatenate(b) if b else np.zeros(0, np.complex64)) for b in bufs]
    for off, p, nsamp, k in tags:
        s = streams[0]
        s.add_tag(off, "cfo", float(p.get("cfo", 0.0)))
        s.add_tag(off, "snr", float(p.get("snr", 0.0)))
        s.add_tag(off, "rssi", float(p.get("rssi", 0.0)))
        s.add_tag(off, "seq", int(p.get("seq", k)))
        s.add_tag(off, "mcs", int(p["l_mcs"]))
        s.add_tag(off, "len", int(p["l_len"]))
        s.add_tag(off, "nsamp", int(nsamp))
        s.add_tag(off, "chan", np.asarray(p["chan"], np.complex64))
    lib().ref_stream_sort_tags(streams[0].h)
    return streams


def make_llr_stream(packets):
    """Build decode-input stream from dicts {'llr': f32[total], 'format','mcs','len','cr','ampdu','trellis', ...}."""
    off = 0
    data = []
    for p in packets:
        data.append(np.asarray(p["llr"], np.float32))
    s = Stream(np.concatenate(data) if data else np.zeros(0, np.float32))
    for p in packets:
        for key in ("cfo", "snr", "rssi", "sssnr0", "sssnr1"):
            if key in p:
                s.add_tag(off, key, float(p[key]))
        for key in ("format", "mcs", "len", "cr", "ampdu", "trellis"):
            s.add_tag(off, key, int(p[key]))
        s.add_tag(off, "total", int(p.get("total", len(p["llr"]))))
        off += len(p["llr"])
    return s


def split_llr_stream(llr_stream):
    """LLR stream from demod -> list of packet dicts (tags + 'llr')."""
    data = llr_stream.data()
    out = []
    for off, t in llr_stream.tags():
        d = dict(t)
        d["offset"] = off
        d["llr"] = data[off:off + d["total"]].copy()
        out.append(d)
    return out


# ---------------------------------------------------------- phy functions
def parse_legacy(mcs, length):
    o = np.zeros(16, np.int32)
    lib().ref_parse_legacy(mcs, length, o.ctypes.data)
    return o


def qam_to_llr(mod, qam):
    q = np.ascontiguousarray(qam, np.complex64)
    nbits = {0: 1, 2: 2, 3: 4, 4: 6, 5: 8}[mod]
    out = np.zeros(q.size * nbits, np.float32)
    lib().ref_qam_to_llr(mod, q.size, q.ctypes.data, out.ctypes.data)
    return out


def deint(which, x):
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros_like(x)
    lib().ref_deint(which, x.size, x.ctypes.data, out.ctypes.data)
    return out


def depas2(nbpscs, a, b):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    out = np.zeros(a.size * 2, np.float32)
    lib().ref_depas2(nbpscs, a.size, a.ctypes.data, b.ctypes.data, out.ctypes.data)
    return out


def sv_decode(llr_pairs, trellis):
    x = np.ascontiguousarray(llr_pairs, np.float32)
    bits = np.zeros(trellis, np.uint8)
    lib().ref_sv_decode(x.ctypes.data, bits.ctypes.data, trellis)
    return bits


def bcc_encode(bits):
    b = np.ascontiguousarray(bits, np.uint8)
    out = np.zeros(b.size * 2, np.uint8)
    lib().ref_bcc_encode(b.ctypes.data, out.ctypes.data, b.size)
    return out


def sv_tables():
    nx = np.zeros(128, np.int32)
    op = np.zeros(128, np.int32)
    lib().ref_sv_tables(nx.ctypes.data, op.ctypes.data)
    return nx.reshape(64, 2), op.reshape(64, 2)


def table(name):
    """One of the reference's constant tables (cloud80211phy.cc) as a float64 array."""
    out = np.zeros(512, np.float64)
    n = lib().ref_table(name.encode(), out.ctypes.data, out.size)
    if n == 0:
        raise KeyError(name)
    return out[:n].copy()


def intel_vhtb20(bits52):
    b = np.ascontiguousarray(bits52, np.uint8)
    out = np.zeros(52, np.uint8)
    lib().ref_intel_vhtb20(b.ctypes.data, out.ctypes.data)
    return out


def check_crc8(bits42):
    b = np.ascontiguousarray(bits42, np.uint8)
    return bool(lib().ref_check_crc8(b.ctypes.data))


def nlsig_demod(sym1, sym2, h):
    a = [np.ascontiguousarray(x, np.complex64) for x in (sym1, sym2, h)]
    ht = np.zeros(96, np.float32)
    vht = np.zeros(96, np.float32)
    lib().ref_nlsig_demod(a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, ht.ctypes.data, vht.ctypes.data)
    return ht, vht
