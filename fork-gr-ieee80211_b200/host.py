"""ctypes binding of libw80211rx.so (include/w80211rx.h) plus a small convenience layer.

Used by tests/, bench.py and __graft_entry__.py.  Every call goes through the
C ABI; there is no Python or CPU implementation behind it: if the shared
library is missing or no CUDA device is usable, `load()` / `Receiver()` raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libw80211rx.so")  # the product library; bench.py --resident-only --lib points tuning runs at a variant build

MEM_HOST, MEM_DEVICE, MEM_LEASE = 0, 1, 2
ST_OK, ST_DROP_GATE, ST_NDP, ST_UNSUPPORTED, ST_BAD_INPUT = 0, 1, 2, 3, 4
F_L, F_HT, F_VHT = 0, 1, 2

SYMBOLS = [
    "w80211_abi_version", "w80211_sizeof", "w80211_create", "w80211_destroy", "w80211_last_error", "w80211_stats_get",
    "w80211_timing_get", "w80211_stream", "w80211_demod_batch", "w80211_decode_batch", "w80211_rx_submit",
    "w80211_rx_collect", "w80211_rx_batch", "w80211_host_alloc", "w80211_host_free", "w80211_signal_batch",
    "w80211_detect_batch", "w80211_presiso_batch", "w80211_viterbi_noalloc", "w80211_rx_poll", "w80211_demod_submit",
    "w80211_decode_submit", "w80211_llr_export", "w80211_llr_query", "w80211_lease_consumers", "w80211_detect_stream",
]


class Cfg(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_packets", C.c_int32), ("nslots", C.c_int32), ("flags", C.c_int32)]


class PktIn(C.Structure):
    _fields_ = [("iq_offset", C.c_uint64), ("nsamp", C.c_int32), ("l_mcs", C.c_int32), ("l_len", C.c_int32),
                ("cfo", C.c_float), ("snr", C.c_float), ("rssi", C.c_float), ("iq_scale", C.c_float)]


PKT_IN_DTYPE = np.dtype([("iq_offset", "<u8"), ("nsamp", "<i4"), ("l_mcs", "<i4"), ("l_len", "<i4"),
                         ("cfo", "<f4"), ("snr", "<f4"), ("rssi", "<f4"), ("iq_scale", "<f4")], align=True)
IQ_FC32, IQ_SC16, IQ_SC8 = 0, 1, 2


class BatchIn(C.Structure):
    _fields_ = [("npkt", C.c_int32), ("nant", C.c_int32), ("mem", C.c_int32), ("iq_format", C.c_int32),
                ("iq", C.c_void_p * 2), ("nsamples", C.c_uint64), ("pkts", C.c_void_p), ("chan", C.c_void_p),
                ("mu_pos", C.c_int32), ("reserved", C.c_int32)]


DESC_DTYPE = np.dtype([
    ("status", "<i4"), ("format", "<i4"), ("mcs", "<i4"), ("len", "<i4"), ("cr", "<i4"), ("ampdu", "<i4"),
    ("trellis", "<i4"), ("total", "<i4"), ("nss", "<i4"), ("nltf", "<i4"), ("nsym", "<i4"), ("nsymsamp", "<i4"),
    ("mod", "<i4"), ("nbpscs", "<i4"), ("ncbpss", "<i4"), ("ncbps", "<i4"), ("ndbps", "<i4"), ("data_off", "<i4"),
    ("sssnr0", "<f4"), ("sssnr1", "<f4"), ("llr_off", "<i8"), ("cfo", "<f4"), ("snr", "<f4"), ("rssi", "<f4"),
    ("l_mcs", "<i4")], align=True)

MPDU_DTYPE = np.dtype([("pkt", "<u4"), ("off", "<u4"), ("len", "<u4"), ("crc_ok", "<u4")])


class BatchOut(C.Structure):
    _fields_ = [("mem", C.c_int32), ("want_llr", C.c_int32), ("desc", C.c_void_p), ("bytes", C.c_void_p),
                ("bytes_capacity", C.c_uint64), ("psdu_off", C.c_void_p), ("mpdus", C.c_void_p),
                ("mpdu_capacity", C.c_uint32), ("llr", C.c_void_p), ("llr_capacity", C.c_uint64),
                ("n_mpdu", C.c_uint32), ("n_crc_ok", C.c_uint32), ("n_bytes", C.c_uint64), ("n_llr", C.c_uint64),
                ("psdu_bits_ok", C.c_uint64)]


class DecodeIn(C.Structure):
    _fields_ = [("npkt", C.c_int32), ("llr_mem", C.c_int32), ("desc", C.c_void_p), ("llr", C.c_void_p), ("nllr", C.c_uint64),
                ("lease", C.c_uint64)]


SYNC_TAGGED = 1
SYNC_DTYPE = np.dtype([("pos", "<u8"), ("rad", "<f4"), ("snr", "<f4"), ("rssi", "<f4"), ("flags", "<i4")], align=True)


class SignalIn(C.Structure):
    _fields_ = [("nsync", C.c_int32), ("nant", C.c_int32), ("mem", C.c_int32), ("final", C.c_int32),
                ("iq", C.c_void_p * 2), ("nsamples", C.c_uint64), ("syncs", C.c_void_p), ("start", C.c_uint64)]


class SignalOut(C.Structure):
    _fields_ = [("mem", C.c_int32), ("npkt", C.c_int32), ("nsync_used", C.c_int32), ("n_rejected", C.c_int32),
                ("next", C.c_uint64), ("need", C.c_uint64), ("iq_capacity", C.c_uint64), ("iq_used", C.c_uint64), ("pkts", C.c_void_p),
                ("sync_of_pkt", C.c_void_p), ("chan", C.c_void_p), ("iq", C.c_void_p * 2),
                ("lsig_ms", C.c_float), ("copy_ms", C.c_float)]


class DetectState(C.Structure):
    _fields_ = [("n_plateau", C.c_int32), ("f_plateau_end", C.c_int32), ("count_down", C.c_int32), ("conj_ac", C.c_float),
                ("conj_re", C.c_float), ("conj_im", C.c_float), ("sync_skip", C.c_int32), ("reserved", C.c_int32)]


class DetectIn(C.Structure):
    _fields_ = [("mem", C.c_int32), ("final", C.c_int32), ("nsamples", C.c_uint64), ("preac", C.c_void_p),
                ("preconj", C.c_void_p), ("sig", C.c_void_p), ("state", C.POINTER(DetectState))]


class DetectOut(C.Structure):
    _fields_ = [("nsync", C.c_int32), ("sync_capacity", C.c_int32), ("n_trigger", C.c_int32), ("reserved", C.c_int32),
                ("next", C.c_uint64), ("syncs", C.c_void_p), ("state", DetectState)]


NOALLOC_MAX_PKTS = 16
NOALLOC_PKT_DTYPE = np.dtype([("llr_off", "<u8"), ("bits_off", "<u8"), ("trellis", "<i4"), ("cr", "<i4")], align=True)


class Stats(C.Structure):
    _fields_ = [("packets_in", C.c_uint64), ("packets_decoded", C.c_uint64), ("dropped_gate", C.c_uint64),
                ("ndp", C.c_uint64), ("unsupported", C.c_uint64), ("bad_input", C.c_uint64),
                ("mpdu_total", C.c_uint64), ("mpdu_crc_ok", C.c_uint64), ("ok_legacy", C.c_uint64 * 8),
                ("ok_ht", C.c_uint64 * 8), ("ok_vht", C.c_uint64 * 10), ("trellis_steps", C.c_uint64),
                ("kernel_launches", C.c_uint64)]


class Timing(C.Structure):
    _fields_ = [("h2d", C.c_float), ("header", C.c_float), ("demod", C.c_float), ("viterbi", C.c_float),
                ("traceback", C.c_float), ("post", C.c_float), ("d2h", C.c_float), ("total", C.c_float),
                ("launches", C.c_uint32), ("host_plan", C.c_float), ("host_collect", C.c_float)]


_lib = None


def load():
    """dlopen the library (no CUDA call yet)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libw80211rx.so is not built: run __graft_entry__.build() (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.w80211_abi_version.restype = C.c_int
        L.w80211_sizeof.restype = C.c_size_t
        L.w80211_sizeof.argtypes = [C.c_int]
        L.w80211_create.argtypes = [C.POINTER(Cfg), C.POINTER(vp)]
        L.w80211_destroy.argtypes = [vp]
        L.w80211_destroy.restype = None
        L.w80211_host_alloc.argtypes = [C.c_size_t]
        L.w80211_host_alloc.restype = vp
        L.w80211_host_free.argtypes = [vp]
        L.w80211_host_free.restype = None
        L.w80211_last_error.argtypes = [vp]
        L.w80211_last_error.restype = C.c_char_p
        L.w80211_stats_get.argtypes = [vp, C.POINTER(Stats)]
        L.w80211_timing_get.argtypes = [vp, C.c_int, C.POINTER(Timing)]
        L.w80211_stream.argtypes = [vp, C.c_int]
        L.w80211_stream.restype = vp
        L.w80211_demod_batch.argtypes = [vp, C.POINTER(BatchIn), C.POINTER(BatchOut)]
        L.w80211_decode_batch.argtypes = [vp, C.c_int32, vp, vp, C.c_uint64, C.c_int32, C.POINTER(BatchOut)]
        L.w80211_rx_submit.argtypes = [vp, C.c_int, C.POINTER(BatchIn), C.POINTER(BatchOut)]
        L.w80211_rx_collect.argtypes = [vp, C.c_int]
        L.w80211_rx_batch.argtypes = [vp, C.POINTER(BatchIn), C.POINTER(BatchOut)]
        L.w80211_signal_batch.argtypes = [vp, C.POINTER(SignalIn), C.POINTER(SignalOut)]
        L.w80211_detect_batch.argtypes = [vp, C.POINTER(DetectIn), C.POINTER(DetectOut)]
        L.w80211_presiso_batch.argtypes = [vp, vp, C.c_uint64, C.c_int32, vp, vp]
        L.w80211_viterbi_noalloc.argtypes = [C.c_int32, vp, C.c_int32, vp, vp, C.c_uint64, vp, C.c_uint64, C.c_char_p, C.c_size_t]
        L.w80211_rx_poll.argtypes = [vp, C.c_int]
        L.w80211_demod_submit.argtypes = [vp, C.c_int, C.POINTER(BatchIn), C.POINTER(BatchOut)]
        L.w80211_decode_submit.argtypes = [vp, C.c_int, C.POINTER(DecodeIn), C.POINTER(BatchOut)]
        L.w80211_llr_export.argtypes = [vp, C.c_int, C.POINTER(C.c_uint64)]
        L.w80211_llr_query.argtypes = [C.c_uint64, C.POINTER(C.c_int32), C.POINTER(C.c_uint64)]
        L.w80211_lease_consumers.argtypes = [C.c_int]
        L.w80211_detect_stream.argtypes = [vp, vp, C.c_uint64, C.c_uint64, C.c_int32, C.c_int32, C.POINTER(DetectState), C.POINTER(DetectOut)]
        _lib = L
    return _lib


class W80211Error(RuntimeError):
    pass


def pack_batch(batches):
    """List of gen_ppdu.gen_batch() dicts (same antenna count) -> (iq list, pkts, chan) numpy arrays:
    the packets' post-L-SIG samples back to back per antenna, the packet table and the legacy H."""
    nant = batches[0]["nrx"]
    iq = [[] for _ in range(nant)]
    pk, ch = [], []
    off = 0
    for b in batches:
        assert b["nrx"] == nant
        npkt, nsamp = b["post"][0].shape
        for a in range(nant):
            iq[a].append(np.ascontiguousarray(b["post"][a], np.complex64).reshape(-1))
        t = np.zeros(npkt, PKT_IN_DTYPE)
        t["iq_offset"] = off + np.arange(npkt, dtype=np.uint64) * nsamp
        t["nsamp"] = nsamp
        t["l_mcs"] = b["l_mcs"]
        t["l_len"] = b["l_len"]
        pk.append(t)
        ch.append(np.ascontiguousarray(b["chan"], np.complex64))
        off += npkt * nsamp
    return [np.concatenate(x) for x in iq], np.concatenate(pk), np.concatenate(ch)


def quantize_sc16(iq, pkts, headroom=0.9, bits=16):
    """fc32 packets -> the sc16 (bits=16) or sc8 (bits=8) wire format: per packet,
    scale = max |component| / (full_scale * headroom), sample = round(x / scale).  Returns (integer arrays [nsamples, 2],
    packet table with iq_scale set, the complex64 arrays a receiver of those integers reconstructs:
    (float)int * scale -- the reference's input for parity)."""
    pk = pkts.copy()
    it, full = (np.int16, 32767.0) if bits == 16 else (np.int8, 127.0)
    q = [np.zeros((len(x), 2), it) for x in iq]
    back = [np.zeros(len(x), np.complex64) for x in iq]
    for p in range(len(pk)):
        o, n = int(pk["iq_offset"][p]), int(pk["nsamp"][p])
        peak = max(float(np.abs(x[o:o + n].view(np.float32)).max()) if n else 0.0 for x in iq)
        scale = np.float32(peak / (full * headroom)) if peak > 0 else np.float32(1.0)
        pk["iq_scale"][p] = scale
        for a, x in enumerate(iq):
            v = np.rint(x[o:o + n].view(np.float32).astype(np.float64) / float(scale)).astype(it).reshape(-1, 2)
            q[a][o:o + n] = v
            back[a][o:o + n] = (v.astype(np.float32) * scale).reshape(-1).view(np.complex64)
    return q, pk, back


class Result:
    """Host-side view of one finished batch."""

    def __init__(self, desc, data, psdu_off, mpdus, out, llr=None):
        self.desc, self.bytes, self.psdu_off, self.mpdus, self.llr = desc, data, psdu_off, mpdus, llr
        self.n_mpdu, self.n_crc_ok, self.psdu_bits_ok = out.n_mpdu, out.n_crc_ok, out.psdu_bits_ok
        self.n_llr, self.n_bytes = out.n_llr, out.n_bytes

    def messages(self):
        """The blobs the reference's decode block publishes (decode_impl.cc:439-441, 512-516), in order:
        [format, len & 0xff, len >> 8, MPDU bytes, mcs] for every FCS-valid MPDU."""
        out = []
        for m in self.mpdus[: self.n_mpdu]:
            if not m["crc_ok"]:
                continue
            d = self.desc[m["pkt"]]
            ln = int(m["len"])
            body = self.bytes[m["off"]: m["off"] + ln].tobytes()
            out.append(bytes([int(d["format"]), ln & 0xFF, ln >> 8]) + body + bytes([int(d["mcs"]) & 0xFF]))
        return out

    def packet_llr(self, p):
        d = self.desc[p]
        return self.llr[d["llr_off"]: d["llr_off"] + d["total"]]


class Receiver:
    """One w80211_ctx (one GPU)."""

    def __init__(self, device=0, nslots=2, viterbi="auto"):
        """viterbi: "auto" | "thread" | "warp" -- which Viterbi kernels decode (w80211_cfg.flags, W80211_CFG_VITERBI_*)."""
        self.L = load()
        self.h = C.c_void_p()
        cfg = Cfg(device, 0, nslots, {"auto": 0, "thread": 1, "warp": 2}[viterbi])
        r = self.L.w80211_create(C.byref(cfg), C.byref(self.h))
        if r != 0:
            raise W80211Error("w80211_create failed (%d): no usable CUDA device; this library has no CPU path" % r)
        self._keep = {}
        self._pinned = []  # page-locked result buffers handed out by pinned_zeros()

    def close(self):
        if self.h:
            self.L.w80211_destroy(self.h)
            self.h = C.c_void_p()
        for p in getattr(self, "_pinned", []):
            self.L.w80211_host_free(p)
        self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, r):
        if r != 0:
            raise W80211Error("%d: %s" % (r, self.L.w80211_last_error(self.h).decode()))

    # -- raw struct helpers -------------------------------------------------
    @staticmethod
    def make_in(iq, pkts, chan, mem=MEM_HOST, nsamples=None, npkt=None, iq_format=None, mu_pos=0):
        """iq: list of numpy arrays (host: complex64, or int16 of shape [nsamples, 2] for sc16) or integer device pointers."""
        bi = BatchIn()
        bi.nant = len(iq)
        bi.mem = mem
        if mem == MEM_HOST:
            bi.npkt = len(pkts)
            bi.nsamples = len(iq[0])
            if iq_format is None:
                iq_format = IQ_SC16 if iq[0].dtype == np.int16 else (IQ_SC8 if iq[0].dtype == np.int8 else IQ_FC32)
            for a, x in enumerate(iq):
                assert x.flags.c_contiguous
                want = {IQ_FC32: np.complex64, IQ_SC16: np.int16, IQ_SC8: np.int8}[iq_format]
                assert x.dtype == want and (iq_format == IQ_FC32 or (x.ndim == 2 and x.shape[1] == 2))
                bi.iq[a] = x.ctypes.data
            assert pkts.dtype == PKT_IN_DTYPE
            bi.pkts = pkts.ctypes.data
            assert chan.dtype == np.complex64 and chan.flags.c_contiguous
            bi.chan = chan.ctypes.data
        else:
            bi.npkt = npkt
            bi.nsamples = nsamples
            for a, x in enumerate(iq):
                bi.iq[a] = int(x)
            bi.pkts = int(pkts)
            bi.chan = int(chan)
        bi.iq_format = int(iq_format or 0)
        bi.mu_pos = int(mu_pos)
        return bi

    def pinned_zeros(self, n, dtype):
        """n zeroed elements in page-locked host memory (w80211_host_alloc), freed by close().  With pageable result buffers
        the device -> host copies at the end of a batch make submit() wait for the whole batch; with these it returns at once."""
        dtype = np.dtype(dtype)
        nbytes = max(int(n) * dtype.itemsize, 16)
        p = self.L.w80211_host_alloc(nbytes)
        if not p:
            raise MemoryError("w80211_host_alloc(%d)" % nbytes)
        self._pinned.append(p)
        a = np.frombuffer((C.c_char * nbytes).from_address(p), dtype=np.uint8)[: int(n) * dtype.itemsize].view(dtype)
        a[...] = np.zeros((), dtype)
        return a

    @staticmethod
    def make_out_host(npkt, bytes_cap, mpdu_cap, llr_cap=0, zeros=np.zeros):
        bo = BatchOut()
        arrs = {
            "desc": zeros(npkt, DESC_DTYPE),
            "bytes": zeros(max(bytes_cap, 16), np.uint8),
            "psdu_off": zeros(max(npkt, 1), np.uint32),
            "mpdus": zeros(max(mpdu_cap, 1), MPDU_DTYPE),
            "llr": zeros(max(llr_cap, 4), np.float32) if llr_cap else None,
        }
        bo.mem = MEM_HOST
        bo.desc = arrs["desc"].ctypes.data
        bo.bytes = arrs["bytes"].ctypes.data
        bo.bytes_capacity = arrs["bytes"].size
        bo.psdu_off = arrs["psdu_off"].ctypes.data
        bo.mpdus = arrs["mpdus"].ctypes.data
        bo.mpdu_capacity = arrs["mpdus"].size
        if llr_cap:
            bo.want_llr = 1
            bo.llr = arrs["llr"].ctypes.data
            bo.llr_capacity = arrs["llr"].size
        return bo, arrs

    # -- convenience --------------------------------------------------------
    def _caps(self, iq, pkts):
        nsamp_tot = int(pkts["nsamp"].astype(np.int64).sum()) if len(pkts) else 0
        # decoded bytes < 1 per sample (2 streams of 256-QAM 5/6: 86 B per 80 samples); LLR floats <= 832 per 72 samples
        return nsamp_tot * 2 + 64 * len(pkts) + 64, len(pkts) * 2 + nsamp_tot // 64 + 64, \
            nsamp_tot * 12 + 64 * len(pkts) + 64

    def rx(self, iq, pkts, chan, want_llr=False, mu_pos=0):
        """IQ -> bytes, fused (w80211_rx_batch)."""
        bcap, mcap, lcap = self._caps(iq, pkts)
        bi = self.make_in(iq, pkts, chan, mu_pos=mu_pos)
        bo, arrs = self.make_out_host(len(pkts), bcap, mcap, lcap if want_llr else 0)
        self._check(self.L.w80211_rx_batch(self.h, C.byref(bi), C.byref(bo)))
        return Result(arrs["desc"], arrs["bytes"], arrs["psdu_off"], arrs["mpdus"], bo, arrs["llr"])

    def submit(self, slot, iq, pkts, chan, want_llr=False, mu_pos=0, bytes_cap=None, pinned=False):
        """IQ -> bytes on slot `slot` without waiting (w80211_rx_submit); collect(slot) returns the Result.
        pinned: result buffers in page-locked memory (see pinned_zeros)."""
        bcap, mcap, lcap = self._caps(iq, pkts)
        bi = self.make_in(iq, pkts, chan, mu_pos=mu_pos)
        bo, arrs = self.make_out_host(len(pkts), bcap if bytes_cap is None else bytes_cap, mcap, lcap if want_llr else 0,
                                      zeros=self.pinned_zeros if pinned else np.zeros)
        keep = (bi, bo, arrs, iq, pkts, chan)  # the library reads them until collect()
        self._check(self.L.w80211_rx_submit(self.h, slot, C.byref(bi), C.byref(bo)))
        self._keep[slot] = keep

    def collect(self, slot):
        try:
            self._check(self.L.w80211_rx_collect(self.h, slot))
            bi, bo, arrs = self._keep[slot][:3]
            return Result(arrs["desc"], arrs["bytes"], arrs["psdu_off"], arrs["mpdus"], bo, arrs["llr"])
        finally:
            self._keep.pop(slot, None)

    def demod(self, iq, pkts, chan, mu_pos=0):
        """IQ -> LLR stream + descriptors (w80211_demod_batch): what demod / demod2 output."""
        bcap, mcap, lcap = self._caps(iq, pkts)
        bi = self.make_in(iq, pkts, chan, mu_pos=mu_pos)
        bo, arrs = self.make_out_host(len(pkts), 16, 1, lcap)
        self._check(self.L.w80211_demod_batch(self.h, C.byref(bi), C.byref(bo)))
        return Result(arrs["desc"], arrs["bytes"], arrs["psdu_off"], arrs["mpdus"], bo, arrs["llr"])

    def decode(self, desc, llr):
        """LLR stream + descriptors -> bytes (w80211_decode_batch): what decode does."""
        desc = np.ascontiguousarray(desc, DESC_DTYPE)
        llr = np.ascontiguousarray(llr, np.float32)
        npkt = len(desc)
        bcap = int(np.maximum(desc["trellis"], 0).astype(np.int64).sum()) // 8 + 64 * npkt + 64
        bo, arrs = self.make_out_host(npkt, bcap, npkt * 2 + bcap // 32 + 64)
        self._check(self.L.w80211_decode_batch(self.h, npkt, desc.ctypes.data, llr.ctypes.data, llr.size, MEM_HOST,
                                               C.byref(bo)))
        return Result(arrs["desc"], arrs["bytes"], arrs["psdu_off"], arrs["mpdus"], bo)

    def demod_submit(self, slot, iq, pkts, chan, want_llr=True, mu_pos=0):
        """IQ -> LLRs + descriptors on slot `slot` without waiting (w80211_demod_submit); collect(slot) returns the Result."""
        bcap, mcap, lcap = self._caps(iq, pkts)
        bi = self.make_in(iq, pkts, chan, mu_pos=mu_pos)
        bo, arrs = self.make_out_host(len(pkts), 16, 1, lcap if want_llr else 0)
        keep = (bi, bo, arrs, iq, pkts, chan)
        self._check(self.L.w80211_demod_submit(self.h, slot, C.byref(bi), C.byref(bo)))
        self._keep[slot] = keep

    def poll(self, slot):
        r = self.L.w80211_rx_poll(self.h, slot)
        if r < 0:
            self._check(r)
        return bool(r)

    def llr_export(self, slot):
        """Detach the LLRs of the demod batch last collected from `slot` as a lease (w80211_llr_export); returns the handle."""
        h = C.c_uint64(0)
        self._check(self.L.w80211_llr_export(self.h, slot, C.byref(h)))
        return int(h.value)

    def llr_query(self, handle):
        dev, n = C.c_int32(-1), C.c_uint64(0)
        r = self.L.w80211_llr_query(int(handle), C.byref(dev), C.byref(n))
        return None if r != 0 else (int(dev.value), int(n.value))

    def decode_submit(self, slot, desc, llr=None, lease=0, pinned=False):
        """Descriptors + LLRs (host array, or a lease handle) -> bytes on slot `slot` without waiting (w80211_decode_submit).
        pinned: result buffers in page-locked memory (see pinned_zeros)."""
        desc = np.ascontiguousarray(desc, DESC_DTYPE)
        npkt = len(desc)
        di = DecodeIn()
        di.npkt, di.desc = npkt, desc.ctypes.data
        if lease:
            di.llr_mem, di.lease = MEM_LEASE, int(lease)
        else:
            llr = np.ascontiguousarray(llr, np.float32)
            di.llr_mem, di.llr, di.nllr = MEM_HOST, llr.ctypes.data, llr.size
        bcap = int(np.maximum(desc["trellis"], 0).astype(np.int64).sum()) // 8 + 64 * npkt + 64
        bo, arrs = self.make_out_host(npkt, bcap, npkt * 2 + bcap // 32 + 64, zeros=self.pinned_zeros if pinned else np.zeros)
        arrs["desc"][:] = desc
        keep = (di, bo, arrs, desc, llr)
        self._check(self.L.w80211_decode_submit(self.h, slot, C.byref(di), C.byref(bo)))
        self._keep[slot] = keep

    def signal(self, raw, pos, rad=0.0, snr=0.0, rssi=0.0, start=0, final=True, tagged=None):
        """Raw samples + sync flags -> packets in demod's input format (w80211_signal_batch): what signal / signal2 do.
        raw: list of 1 or 2 complex64 arrays; pos: flag positions (ascending); rad/snr/rssi: scalars or per-flag arrays.
        Returns dict(iq=[per antenna], pkts, chan, sync_of_pkt, next, nsync_used, n_rejected)."""
        raw = [np.ascontiguousarray(x, np.complex64) for x in raw]
        pos = np.atleast_1d(np.asarray(pos, np.uint64))
        sy = np.zeros(len(pos), SYNC_DTYPE)
        sy["pos"] = pos
        sy["rad"] = rad
        sy["snr"] = snr
        sy["rssi"] = rssi
        sy["flags"] = SYNC_TAGGED if tagged is None else np.where(np.asarray(tagged), SYNC_TAGGED, 0)
        n = len(raw[0])
        si = SignalIn()
        si.nsync, si.nant, si.mem, si.final = len(sy), len(raw), MEM_HOST, int(bool(final))
        for a, x in enumerate(raw):
            assert len(x) == n
            si.iq[a] = x.ctypes.data
        si.nsamples = n
        si.syncs = sy.ctypes.data
        si.start = int(start)
        so = SignalOut()
        cap = max(len(sy), 1)
        pk = np.zeros(cap, PKT_IN_DTYPE)
        sop = np.zeros(cap, np.int32)
        ch = np.zeros((cap, 64), np.complex64)
        out = [np.zeros(max(n, 1), np.complex64) for _ in raw]
        so.mem = MEM_HOST
        so.iq_capacity = n
        so.pkts, so.sync_of_pkt, so.chan = pk.ctypes.data, sop.ctypes.data, ch.ctypes.data
        for a, x in enumerate(out):
            so.iq[a] = x.ctypes.data
        self._check(self.L.w80211_signal_batch(self.h, C.byref(si), C.byref(so)))
        k, used = so.npkt, so.iq_used
        return {"iq": [x[:used] for x in out], "pkts": pk[:k].copy(), "chan": ch[:k].copy(), "sync_of_pkt": sop[:k].copy(),
                "next": int(so.next), "need": int(so.need), "nsync_used": int(so.nsync_used), "n_rejected": int(so.n_rejected)}

    def presiso(self, sig):
        """samples -> (preac, preconj) on the GPU (w80211_presiso_batch): examples/presiso.grc."""
        sig = np.ascontiguousarray(sig, np.complex64)
        ac = np.zeros(len(sig), np.float32)
        cj = np.zeros(len(sig), np.complex64)
        self._check(self.L.w80211_presiso_batch(self.h, sig.ctypes.data, len(sig), MEM_HOST, ac.ctypes.data, cj.ctypes.data))
        return ac, cj

    def detect(self, preac, preconj, sig, state=None, final=True):
        """presiso streams -> sync flags (w80211_detect_batch): what trigger + sync do.
        Returns dict(syncs=SYNC_DTYPE array, next, state=DetectState, n_trigger)."""
        preac = np.ascontiguousarray(preac, np.float32)
        preconj = np.ascontiguousarray(preconj, np.complex64)
        sig = np.ascontiguousarray(sig, np.complex64)
        n = len(preac)
        assert len(preconj) == n and len(sig) == n
        di, do = DetectIn(), DetectOut()
        di.mem, di.final, di.nsamples = MEM_HOST, int(bool(final)), n
        di.preac, di.preconj, di.sig = preac.ctypes.data, preconj.ctypes.data, sig.ctypes.data
        if state is not None:
            di.state = C.pointer(state)
        cap = n // 100 + 16
        sy = np.zeros(cap, SYNC_DTYPE)
        do.sync_capacity, do.syncs = cap, sy.ctypes.data
        self._check(self.L.w80211_detect_batch(self.h, C.byref(di), C.byref(do)))
        st = DetectState()
        C.memmove(C.byref(st), C.byref(do.state), C.sizeof(DetectState))
        return {"syncs": sy[:do.nsync].copy(), "next": int(do.next), "state": st, "n_trigger": int(do.n_trigger)}

    def detect_stream(self, sig, lead=0, state=None, final=True):
        """samples (the first `lead` are history) -> sync flags, presiso streams kept on the GPU (w80211_detect_stream)."""
        sig = np.ascontiguousarray(sig, np.complex64)
        n = len(sig)
        do = DetectOut()
        cap = n // 100 + 16
        sy = np.zeros(cap, SYNC_DTYPE)
        do.sync_capacity, do.syncs = cap, sy.ctypes.data
        self._check(self.L.w80211_detect_stream(self.h, sig.ctypes.data, n, int(lead), MEM_HOST, int(bool(final)),
                                                C.pointer(state) if state is not None else None, C.byref(do)))
        st = DetectState()
        C.memmove(C.byref(st), C.byref(do.state), C.sizeof(DetectState))
        return {"syncs": sy[:do.nsync].copy(), "next": int(do.next), "state": st, "n_trigger": int(do.n_trigger)}

    def stats(self):
        s = Stats()
        self._check(self.L.w80211_stats_get(self.h, C.byref(s)))
        return s

    def timing(self, slot=0):
        t = Timing()
        self._check(self.L.w80211_timing_get(self.h, slot, C.byref(t)))
        return t
