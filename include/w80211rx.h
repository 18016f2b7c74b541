/* w80211rx.h -- C ABI of libw80211rx.so: the B200 (sm_100a) implementation of the
 * gr-ieee80211 receive hot path  demod / demod2  ->  decode.
 *
 * This is the drop-in boundary.  The three GNU Radio blocks of the reference
 *     gr::ieee80211::demod::make(int mupos, int mugid)   include/gnuradio/ieee80211/demod.h:35
 *     gr::ieee80211::demod2::make()                      include/gnuradio/ieee80211/demod2.h:35
 *     gr::ieee80211::decode::make(bool ifdebug)          include/gnuradio/ieee80211/decode.h:35
 * keep their factories, io signatures, tags and messages (gr/ in this repo);
 * their general_work() bodies gather whole packets and call the entry points
 * below instead of the per-symbol / per-trellis-step CPU loops
 *     demod_impl.cc:59-342, demod2_impl.cc:58-349   -> w80211_demod_batch()
 *     decode_impl.cc:60-520                           -> w80211_decode_batch()
 * and a batch receiver that keeps the LLRs on the device calls
 *     w80211_rx_submit() / w80211_rx_collect()        (demod + decode fused, double buffered).
 * The upstream neighbour signal / signal2 (include/gnuradio/ieee80211/signal.h:35, signal2.h:35) is covered by
 *     signal_impl.cc:62-212, signal2_impl.cc:62-222   -> w80211_signal_batch()
 *
 * Conventions: plain C structs, caller owns every host buffer, the library owns
 * device memory and streams; int return, 0 = ok, < 0 = error (w80211_last_error());
 * one ctx per GPU; a ctx is used from one thread at a time (GNU Radio calls a
 * block's general_work from exactly one thread).  There is no CPU fallback:
 * every entry point fails if no CUDA device is usable.
 */
#ifndef W80211RX_H
#define W80211RX_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define W80211_ABI_VERSION 8

/* packet formats: the reference's C8P_F_L / C8P_F_HT / C8P_F_VHT (lib/cloud80211phy.h:35-37) */
#define W80211_F_L 0
#define W80211_F_HT 1
#define W80211_F_VHT 2
/* code rates: C8P_CR_12 .. C8P_CR_56 (lib/cloud80211phy.h:46-49) */
#define W80211_CR_12 0
#define W80211_CR_23 1
#define W80211_CR_34 2
#define W80211_CR_56 3

/* limits of the reference's decode block (lib/decode_impl.h:35-36) */
#define W80211_B_MAX 4095
#define W80211_T_MAX 32782

/* where the caller's sample / LLR / output buffers live */
#define W80211_MEM_HOST 0   /* host memory (pinned or pageable); copies are part of the call */
#define W80211_MEM_DEVICE 1 /* device memory of the ctx's GPU; nothing is copied */
#define W80211_MEM_LEASE 2  /* w80211_decode_in only: the LLRs are a demod batch still in HBM, named by a lease handle */

/* sample format of w80211_batch_in.iq[] (the IQ wire format of the host feed) */
#define W80211_IQ_FC32 0 /* interleaved (re, im) binary32, what UHD hands to the reference's flowgraph */
#define W80211_IQ_SC16 1 /* interleaved (re, im) int16, the native USRP wire format; sample = (float)int16 * pkt.iq_scale,
                            one binary32 multiply per component, so a reference fed the same products sees identical input */

#define W80211_IQ_SC8 2  /* interleaved (re, im) int8 (UHD's otw_format=sc8): sample = (float)int8 * pkt.iq_scale; a quarter of
                            the fc32 bytes at about 39 dB of quantisation SNR when the packet's peak maps to 0.9 full scale */

/* error codes */
#define W80211_OK 0
#define W80211_E_ARG (-1)
#define W80211_E_CUDA (-2)
#define W80211_E_NOMEM (-3)
#define W80211_E_STATE (-4)
#define W80211_E_CAPACITY (-5)

/* Per-packet verdict of the header stage (format detection, SIG parsing, gates).
 * Everything but W80211_ST_OK means "the reference's demod goes to its CLEAN state
 * without emitting LLRs" (demod_impl.cc:159,186) or "outside the scope of this library". */
#define W80211_ST_OK 0
#define W80211_ST_DROP_GATE 1    /* len / nSS / L-SIG-duration gate failed (demod_impl.cc:159,186) */
#define W80211_ST_NDP 2          /* VHT NDP (nSym == 0): demod emits the mu2x1chan tag, no LLRs to decode */
#define W80211_ST_UNSUPPORTED 3  /* VHT MU group id, VHT MCS > 9: not handled */
#define W80211_ST_BAD_INPUT 4    /* descriptor inconsistent (nsamp too short for the header fields, l_mcs out of range) */

typedef struct w80211_ctx w80211_ctx;

typedef struct w80211_cfg {
    int32_t device;        /* CUDA device ordinal */
    int32_t max_packets;   /* capacity hint per batch (grown on demand) */
    int32_t nslots;        /* double buffering depth for rx_submit/collect: 1..4 (0 = 2) */
    int32_t flags;         /* W80211_CFG_* below, 0 = defaults */
} w80211_cfg;

/* w80211_cfg.flags, bits 0-1: which Viterbi kernels decode a code-rate group (both are bit-exact, tests run every parity
 * case under both).  AUTO: one warp per packet (64 path metrics over the 32 lanes, ballot survivors, parallel exact
 * traceback) while a launch has too few packets to give every lane of the GPU a packet of its own, else one thread per
 * packet; the crossover is W80211_WARP_MAX_PKTS packets per code rate (environment, default 7168). */
#define W80211_CFG_VITERBI_AUTO 0
#define W80211_CFG_VITERBI_THREAD 1 /* always one thread per packet (csrc/viterbi.cu viterbi_kernel) */
#define W80211_CFG_VITERBI_WARP 2   /* always one warp per packet (viterbi_warp_kernel) */
#define W80211_CFG_VITERBI_MASK 3

/* One packet as the reference's signal / signal2 block hands it over
 * (tags written at lib/signal_impl.cc:135-152). */
typedef struct w80211_pkt_in {
    uint64_t iq_offset; /* index (in samples) of the packet's first sample after L-SIG in every antenna array */
    int32_t nsamp;      /* tag "nsamp": samples that belong to the packet (without the 320 pad items) */
    int32_t l_mcs;      /* tag "mcs": legacy rate index 0..7 */
    int32_t l_len;      /* tag "len": L-SIG LENGTH */
    float cfo, snr, rssi; /* tags carried through to the output untouched */
    float iq_scale;     /* W80211_IQ_SC16 / SC8 only: value of one integer unit for this packet */
} w80211_pkt_in;

typedef struct w80211_batch_in {
    int32_t npkt;
    int32_t nant;            /* 1: demod (lib/demod_impl.cc), 2: demod2 (lib/demod2_impl.cc) */
    int32_t mem;             /* W80211_MEM_HOST / W80211_MEM_DEVICE: applies to iq[], chan, pkts */
    int32_t iq_format;       /* W80211_IQ_FC32 (0), W80211_IQ_SC16 or W80211_IQ_SC8 */
    const void* iq[2];       /* per antenna: interleaved (re,im) samples, float, int16 or int8 according to iq_format */
    uint64_t nsamples;       /* length of each iq[] array in complex samples */
    const w80211_pkt_in* pkts; /* [npkt] */
    const float* chan;       /* [npkt][64][2]: tag "chan", legacy channel in natural FFT order (signal_impl.cc:143) */
    int32_t mu_pos;          /* nant = 1 only: this receiver's user position (0 or 1) in a VHT MU PPDU, the `mupos` argument of
                                demod::make (demod_impl.cc:34-39, 347-379); SU packets ignore it */
    int32_t reserved;
} w80211_batch_in;

/* Per-packet parameters: the reference's c8p_mod (lib/cloud80211phy.h:58-98) plus the
 * demod->decode tag values (demod_impl.cc:224-253) and where the packet's LLRs are. */
typedef struct w80211_desc {
    int32_t status;   /* W80211_ST_* */
    int32_t format;   /* tag "format" */
    int32_t mcs;      /* tag "mcs" */
    int32_t len;      /* tag "len" */
    int32_t cr;       /* tag "cr" */
    int32_t ampdu;    /* tag "ampdu" */
    int32_t trellis;  /* tag "trellis" */
    int32_t total;    /* tag "total" = nSym * nCBPS */
    int32_t nss, nltf, nsym, nsymsamp;
    int32_t mod, nbpscs, ncbpss, ncbps;
    int32_t ndbps;
    int32_t data_off; /* first data symbol, in samples from iq_offset */
    float sssnr0, sssnr1; /* tags "sssnr0"/"sssnr1" (VHT only) */
    int64_t llr_off;  /* float index of the packet's first LLR in the LLR buffer (multiple of 4) */
    float cfo, snr, rssi;
    int32_t l_mcs;    /* copy of the input rate index */
} w80211_desc;

/* One decoded MPDU (legacy / HT: one per packet; VHT: one per A-MPDU subframe walked). */
typedef struct w80211_mpdu {
    uint32_t pkt;     /* packet index in the batch */
    uint32_t off;     /* byte offset of the MPDU inside out_bytes */
    uint32_t len;     /* MPDU length in bytes (FCS included) */
    uint32_t crc_ok;  /* 1: CRC-32 residue == 558161692 (decode_impl.cc:374,453) -> the reference publishes it */
} w80211_mpdu;

typedef struct w80211_batch_out {
    int32_t mem;             /* where desc / bytes / mpdus / llr live (HOST or DEVICE); counts are always host */
    int32_t want_llr;        /* also return the LLR stream (what demod writes to its output port) */
    w80211_desc* desc;       /* [npkt] */
    uint8_t* bytes;          /* decoded, descrambled PSDU bytes of all packets, packet p at psdu_off[p] */
    uint64_t bytes_capacity;
    uint32_t* psdu_off;      /* [npkt] byte offset of bit 16 (first PSDU byte) of packet p in bytes */
    w80211_mpdu* mpdus;      /* [mpdu_capacity], sorted by (pkt, off); NULL: records are counted but not returned */
    uint32_t mpdu_capacity;  /* one record per legacy / HT packet; a VHT packet yields one per A-MPDU delimiter the reference's walk
                                visits, FCS-valid or not (zero-length ones included): at most trellis / 32 + 1.  A batch with more
                                records than this returns W80211_E_CAPACITY; the library's own buffers grow to the exact count */
    float* llr;              /* [llr_capacity] or NULL */
    uint64_t llr_capacity;
    /* results (host side) */
    uint32_t n_mpdu;         /* entries written to mpdus */
    uint32_t n_crc_ok;       /* of which crc_ok */
    uint64_t n_bytes;        /* bytes used */
    uint64_t n_llr;          /* floats used */
    uint64_t psdu_bits_ok;   /* 8 * sum(len of crc_ok MPDUs): the numerator of the throughput metric */
} w80211_batch_out;

typedef struct w80211_stats {
    uint64_t packets_in, packets_decoded, dropped_gate, ndp, unsupported, bad_input;
    uint64_t mpdu_total, mpdu_crc_ok;
    uint64_t ok_legacy[8], ok_ht[8], ok_vht[10]; /* the counters of decode_impl.h:88-90 */
    uint64_t trellis_steps;  /* sum of trellis lengths decoded */
    uint64_t kernel_launches;
} w80211_stats;

/* ---- the front half: signal / signal2 ------------------------------------------
 * One sync flag of the reference's sync block: a non-zero item on signal's input 0 with the tags
 * "rad", "snr", "rssi" (lib/signal_impl.cc:80-98). */
#define W80211_SYNC_TAGGED 1 /* flags: the item carries tags; a flag without tags is skipped (signal_impl.cc:100-104) */
typedef struct w80211_sync {
    uint64_t pos;    /* index of the flagged sample ("L-LTF start + 16") in iq[] */
    float rad;       /* tag "rad": carrier frequency offset in radians per sample */
    float snr, rssi; /* tags "snr", "rssi": carried through */
    int32_t flags;   /* W80211_SYNC_TAGGED */
} w80211_sync;

typedef struct w80211_signal_in {
    int32_t nsync;
    int32_t nant;            /* 1: signal (lib/signal_impl.cc), 2: signal2 (lib/signal2_impl.cc; L-SIG from antenna 1) */
    int32_t mem;             /* where iq[] lives; syncs is always host memory */
    int32_t final;           /* non-zero: no samples follow this buffer, a packet cut off by its end is dropped */
    const float* iq[2];      /* raw received samples, interleaved (re,im) binary32 */
    uint64_t nsamples;       /* complex samples per antenna */
    const w80211_sync* syncs; /* [nsync], ascending pos */
    uint64_t start;          /* the trigger search begins here (0, or what the previous buffer left over) */
} w80211_signal_in;

typedef struct w80211_signal_out {
    int32_t mem;             /* where iq[] and chan are written: a MEM_DEVICE result feeds w80211_rx_submit directly */
    int32_t npkt;            /* out: packets whose L-SIG passed signalCheckLegacy and that lie wholly inside the buffer */
    int32_t nsync_used;      /* out: leading entries of syncs[] that were settled (accepted, rejected, skipped or overlapped) */
    int32_t n_rejected;      /* out: flags whose L-SIG failed the check (the reference resumes 80 samples later) */
    uint64_t next;           /* out: where the reference's state machine stands: first sample the next call must still see */
    uint64_t need;           /* out: buffer length (from sample 0 of this buffer) at which the flag at `next` can be settled;
                                0 when nothing is pending */
    uint64_t iq_capacity;    /* complex samples per antenna iq[] can hold */
    uint64_t iq_used;        /* out */
    w80211_pkt_in* pkts;     /* host [>= nsync]: iq_offset into iq[], nsamp, l_mcs, l_len, cfo = rad * 20e6 / 2pi, snr, rssi */
    int32_t* sync_of_pkt;    /* host [>= nsync] or NULL: index into syncs[] of each packet */
    float* chan;             /* [>= nsync][64][2]: tag "chan" (signal_impl.cc:143) */
    float* iq[2];            /* CFO-corrected samples after L-SIG, nsamp per packet, packets back to back (no pad items) */
    float lsig_ms, copy_ms;  /* out: CUDA-event durations of the two kernels of this call */
} w80211_signal_out;

/* ---- packet detection: trigger + sync ----------------------------------------------
 * State of the reference's trigger block (lib/trigger_impl.h:38-43) and sync block (lib/sync_impl.h:40-52) between two
 * buffers of one stream.  All zero for a new stream. */
typedef struct w80211_detect_state {
    int32_t n_plateau;     /* trigger: length of the current run of preac > 0.3 (d_nPlateau) */
    int32_t f_plateau_end; /* trigger: this run has already started a countdown (d_fPlateauEnd) */
    int32_t count_down;    /* trigger: samples until the pending trigger flag, 0 = none (d_fPlateau / d_countDown) */
    float conj_ac;         /* trigger: largest preac of the current run (d_conjAc) */
    float conj_re, conj_im; /* sync: d_conjMultiAvg, the preconj sample at the last 0x02 flag it saw */
    int32_t sync_skip;     /* sync: leading samples of the next buffer that its SYNC state has already consumed */
    int32_t reserved;
} w80211_detect_state;

typedef struct w80211_detect_in {
    int32_t mem;           /* where preac / preconj / sig live */
    int32_t final;         /* non-zero: no samples follow; a trigger closer than 240 samples to the end is dropped */
    uint64_t nsamples;
    const float* preac;    /* [nsamples] presiso output 0: |moving average 48 of x[n-16] conj(x[n])| / moving average 64 of |x|^2 */
    const float* preconj;  /* [nsamples][2] presiso output 1: that complex moving average (examples/presiso.grc) */
    const float* sig;      /* [nsamples][2] the received samples, aligned with the two streams above (examples/rx.grc:753-767) */
    const w80211_detect_state* state; /* NULL = new stream */
} w80211_detect_in;

typedef struct w80211_detect_out {
    int32_t nsync;         /* out: sync flags found, = entries written to syncs[] */
    int32_t sync_capacity; /* entries syncs[] can hold */
    int32_t n_trigger;     /* out: trigger flags (0x01) the sync block acted on */
    int32_t reserved;
    uint64_t next;         /* out: first sample the next call must start at (nsamples unless a trigger waits for samples) */
    w80211_sync* syncs;    /* host: position, rad, snr, rssi of every flag, ascending: the input of w80211_signal_batch */
    w80211_detect_state state; /* out: state at `next` */
} w80211_detect_out;

/* Timing of the last completed batch, measured with CUDA events on the ctx's own stream (ms). */
typedef struct w80211_timing {
    float h2d, header, demod, viterbi, traceback, post, d2h, total;
    uint32_t launches; /* kernels launched for that batch */
    float host_plan;    /* host wall time (ms) between the header kernel's results and the rest being enqueued */
    float host_collect; /* host wall time (ms) spent in collect after the stream finished (record sort, counters) */
} w80211_timing;

int w80211_abi_version(void);
/* sizeof of the ABI structs as compiled: 0 pkt_in, 1 batch_in, 2 desc, 3 mpdu, 4 batch_out, 5 stats, 6 timing, 7 cfg,
 * 8 sync, 9 signal_in, 10 signal_out, 11 detect_state, 12 detect_in, 13 detect_out, 14 decode_in
 * (lets a binding written in another language check its own layout) */
size_t w80211_sizeof(int what);
int w80211_create(const w80211_cfg* cfg, w80211_ctx** out);
void w80211_destroy(w80211_ctx* ctx);
const char* w80211_last_error(const w80211_ctx* ctx);
int w80211_stats_get(const w80211_ctx* ctx, w80211_stats* out);
int w80211_timing_get(const w80211_ctx* ctx, int slot, w80211_timing* out);
/* the CUDA stream (cudaStream_t) slot `slot` launches on, for callers that time with their own events */
void* w80211_stream(const w80211_ctx* ctx, int slot);

/* Page-locked host memory for staging buffers, so that a caller (the GNU Radio blocks in gr/lib) gets
 * full-speed DMA without linking the CUDA runtime itself.  NULL when the allocation fails. */
void* w80211_host_alloc(size_t bytes);
void w80211_host_free(void* p);

/* Received samples -> the two presiso streams (examples/presiso.grc: delay 16, multiply conjugate, moving averages of 48
 * and 64, magnitude, divide), sig / preac / preconj all in `mem`.  Samples before sig[0] count as zero: a streaming caller
 * overlaps consecutive buffers by 79 samples or more and ignores the first 79 outputs.  The reference holds no source for
 * this stage (stock GNU Radio blocks with running binary32 sums), so the result is defined to 1e-4 relative against a
 * binary64 evaluation.  Synchronous, on the signal stage's stream. */
int w80211_presiso_batch(w80211_ctx* ctx, const float* sig, uint64_t nsamples, int32_t mem, float* preac, float* preconj);

/* presiso streams -> sync flags: what the trigger block (lib/trigger_impl.cc:60-117) and the sync block
 * (lib/sync_impl.cc:61-196) do between presiso and signal.  The runs of preac > 0.3 are found in parallel; the host
 * walks them the way the trigger's plateau / countdown logic and the sync block's IDLE / SYNC states would; one GPU
 * thread per surviving trigger then repeats sync_impl::ltf_autoCorrelation and ltf_cfo with the reference's own
 * sequential binary32 running sums.  Flag positions and rssi are exact; rad and snr go through atan2f / cosf / sinf /
 * log10 and agree with a host libm to a few ulp.  Synchronous, on the signal stage's stream. */
int w80211_detect_batch(w80211_ctx* ctx, const w80211_detect_in* in, w80211_detect_out* out);

/* Received samples -> sync flags in one call: w80211_presiso_batch + w80211_detect_batch with the two presiso streams kept
 * in HBM (what presiso.grc + trigger + sync do between the sample source and signal; this is what gr/lib's `detect` block
 * calls).  sig holds `lead` samples of history (at most what the previous call left unexamined plus 79 or more before it; 0
 * at the start of a stream, where earlier samples count as zero) followed by the nsamples - lead samples to examine;
 * positions in *out (syncs[].pos, next) count from sig + lead.  state / final / out as for w80211_detect_batch. */
int w80211_detect_stream(w80211_ctx* ctx, const float* sig, uint64_t nsamples, uint64_t lead, int32_t mem, int32_t final,
                         const w80211_detect_state* state, w80211_detect_out* out);

/* Raw samples + sync flags -> packets in demod's input format: what signal / signal2 do between sync and demod
 * (lib/signal_impl.cc:62-212, lib/signal2_impl.cc:62-222).  Every flag's L-SIG is demodulated and decoded in
 * parallel (CFO rotation, 3 FFTs, procLHSigDemodDeint cloud80211phy.cc:609-627, 24-step Viterbi, signalCheckLegacy
 * :650-728); the host then walks the flags in order the way the block's TRIGGER/DEMOD/COPY states would (a flag
 * inside an accepted packet, or within 80 samples after a rejected one, is never looked at) and a second kernel
 * copies the accepted packets with the per-sample rotation of signal_impl.cc:172-176.  Synchronous, on a stream
 * of its own (it may run while rx slots are in flight).
 * The rotation uses cosf/sinf: results agree with a host libm to 1 ulp of the factor, not bit for bit. */
int w80211_signal_batch(w80211_ctx* ctx, const w80211_signal_in* in, w80211_signal_out* out);

/* IQ -> LLR stream + descriptors: what demod / demod2 write to their output port.
 * out->desc and out->llr are filled; nothing is decoded.  Synchronous. */
int w80211_demod_batch(w80211_ctx* ctx, const w80211_batch_in* in, w80211_batch_out* out);

/* LLR stream + descriptors -> PSDU bytes + FCS verdicts: what decode does with demod's output.
 * desc[p].{format,mcs,len,cr,ampdu,trellis,total,llr_off} are read (llr_off may be any float index);
 * llr_mem says where desc and llr live.  Synchronous. */
int w80211_decode_batch(w80211_ctx* ctx, int32_t npkt, const w80211_desc* desc, const float* llr,
                        uint64_t nllr, int32_t llr_mem, w80211_batch_out* out);

/* Fused path: IQ -> bytes, LLRs stay in HBM.  submit() enqueues the input copies and the header stage on the slot's
 * stream and returns without waiting; the rest of the batch (host plan, demod, decode, result copies) is enqueued by
 * the next call of submit() / collect() on this context that finds the header stage finished, in submission order, so a
 * caller that keeps 2-4 slots in flight never blocks on a copy.  *in's arrays and *out must stay valid until collect(),
 * which waits for the slot, reports any error of that second half and fills the host-side counters of *out.
 * Device-resident results (out->mem = W80211_MEM_DEVICE) are written straight into out->bytes when it is 16-byte
 * aligned.  rx_batch() = submit + collect on slot 0. */
int w80211_rx_submit(w80211_ctx* ctx, int slot, const w80211_batch_in* in, w80211_batch_out* out);
int w80211_rx_collect(w80211_ctx* ctx, int slot);
int w80211_rx_batch(w80211_ctx* ctx, const w80211_batch_in* in, w80211_batch_out* out);
/* 1: w80211_rx_collect(ctx, slot) would not wait (the slot's batch has finished on the device), 0: still running, < 0: error.
 * Also enqueues the second half of any batch whose header stage has finished, like every other call on the context. */
int w80211_rx_poll(w80211_ctx* ctx, int slot);

/* ---- the two blocks as separate, non-blocking stages (what gr/lib's demod / demod2 and decode call) ------------------
 * The reference's demod and decode are two blocks joined by a float stream (demod_impl.cc:224-253 -> decode_impl.cc:60-121).
 * demod_submit() is rx_submit() without the decode half: descriptors (and, with out->want_llr, the LLR stream) come back
 * at w80211_rx_collect().  After that collect the batch's LLRs are still in HBM; w80211_llr_export() detaches them from the
 * slot as a LEASE, a process-wide handle another context on the same GPU -- the decode block's -- can decode from without
 * the 43 LLRs per payload byte crossing PCIe twice.  A producer always keeps its W80211_LEASES_KEPT newest leases, and older
 * ones too while it holds at most 64 leases and W80211_LEASE_BYTES of device memory in them (environment, default 4 GiB: a
 * consumer may lag by whatever the edge between the blocks buffers); beyond that an export retires the oldest lease that no
 * decode call is reading (its handle turns invalid, its memory is reused).
 * decode_submit() is the non-blocking form of decode_batch(): LLRs from host memory (in->llr must stay valid until
 * collect), device memory, or a lease (in->llr ignored; desc[p].llr_off are the offsets demod reported).  Results arrive
 * at w80211_rx_collect(ctx, slot) in *out exactly as for decode_batch(). */
#define W80211_LEASES_KEPT 6
typedef struct w80211_decode_in {
    int32_t npkt;
    int32_t llr_mem;            /* W80211_MEM_HOST / W80211_MEM_DEVICE / W80211_MEM_LEASE: where llr (and, for DEVICE, desc) live */
    const w80211_desc* desc;    /* [npkt]; always host memory for MEM_LEASE */
    const float* llr;
    uint64_t nllr;              /* floats in llr (ignored for MEM_LEASE) */
    uint64_t lease;             /* MEM_LEASE: handle from w80211_llr_export() */
} w80211_decode_in;
int w80211_demod_submit(w80211_ctx* ctx, int slot, const w80211_batch_in* in, w80211_batch_out* out);
int w80211_decode_submit(w80211_ctx* ctx, int slot, const w80211_decode_in* in, w80211_batch_out* out);
/* slot must have been collected and not reused; *handle != 0 on success */
int w80211_llr_export(w80211_ctx* ctx, int slot, uint64_t* handle);
/* 0 and the lease's device / length when the handle is still valid, W80211_E_STATE when it has been retired */
int w80211_llr_query(uint64_t handle, int32_t* device, uint64_t* nfloats);
/* Consumers of leases in this process (each decode block adds 1 while it lives): a demod block in its default hand-off
 * mode leaves the LLRs on the device only while this is > 0.  delta = 0 just reads the count. */
int w80211_lease_consumers(int delta);

/* ---- no-allocation decode (no counterpart in the reference) -----------
 * LLRs -> decoded trellis bits (what vstb_update + vstb_end leave in v_scramBits, lib/decode_impl.cc:205-302: still scrambled,
 * bit i of word j = trellis bit 32 j + i) for callers that cannot call cudaMalloc: code under CUDA graph capture, tight
 * serving loops, harnesses that forbid allocation.  The call allocates nothing (no cudaMalloc / cudaFree, no library
 * routine that allocates) and needs no w80211_ctx: llr and bits are the caller's device buffers on `device`, `stream` is the
 * caller's cudaStream_t (NULL = the default stream), pkts is a host array that travels as a kernel argument.  The survivor
 * words of the Viterbi forward pass live in a static device array inside the library, W80211_NOALLOC_MAX_PKTS x 32784 x 8 B
 * = 4.0 MiB, held by every GPU that loads the library until the process exits; hence the limits: at most
 * W80211_NOALLOC_MAX_PKTS packets per call (more: W80211_E_CAPACITY and a message in err), trellis <= W80211_T_MAX.
 * The kernels are enqueued on `stream` and the call returns without waiting.  All calls share the static array, which
 * keeps the last call's contents (nothing in it has to start at zero: every word that is read was written by the same
 * call first): calls that overlap on the GPU -- other streams, other threads -- race; serialise them on one stream.
 * Same arithmetic as w80211_decode_batch, so the bits are identical to what it decodes (tests/test_zz_no_alloc.py). */
#define W80211_NOALLOC_MAX_PKTS 16
typedef struct w80211_noalloc_pkt {
    uint64_t llr_off;  /* index of the packet's first LLR in llr[], a multiple of 4 (llr itself 16-byte aligned) */
    uint64_t bits_off; /* index of the packet's first output word in bits[]; (trellis + 31) / 32 words are written */
    int32_t trellis;   /* tag "trellis", 1 .. W80211_T_MAX */
    int32_t cr;        /* tag "cr", W80211_CR_12 .. W80211_CR_56; the packet's LLRs are punctured accordingly */
} w80211_noalloc_pkt;
/* err (may be NULL): receives a message when the return value is not W80211_OK */
int w80211_viterbi_noalloc(int32_t device, void* stream, int32_t npkt, const w80211_noalloc_pkt* pkts, const float* llr,
                           uint64_t nllr, uint32_t* bits, uint64_t nwords, char* err, size_t errlen);

#ifdef __cplusplus
}
#endif
#endif /* W80211RX_H */
