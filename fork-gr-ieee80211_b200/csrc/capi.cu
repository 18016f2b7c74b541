/* capi.cu -- host side of libw80211rx.so: the C ABI of include/w80211rx.h.
 *
 * One ctx per GPU.  A ctx owns `nslots` independent slots (stream + events + device buffers +
 * pinned staging); a slot runs one batch at a time:
 *
 *   H2D(IQ, packet table, legacy H)            [host memory input only]
 *   header kernel   -> descriptors             hdr.cu
 *   D2H(descriptors), host plan                LLR offsets, symbol prefix sums, packets sorted by
 *                                              (code rate, trellis length) into groups of 32, output offsets
 *   H2D(plan)
 *   demod kernel    -> LLRs (stay in HBM)      demod.cu
 *   viterbi kernel  -> survivor words          viterbi.cu
 *   traceback kernel-> decoded bit words       viterbi.cu
 *   post kernel     -> PSDU bytes, MPDU records, counters   post.cu
 *   D2H(bytes, records, counters)
 *
 * The host plan is the only host step inside a batch, and no call waits for it: rx_submit() returns after enqueueing
 * the header stage, and the second half of the batch is enqueued by the next API call that finds the descriptors on the
 * host (progress()).  With several slots in flight a batch's demod kernel waits for the previous batch's traceback and
 * post kernels, so the slots alternate and one slot's host work hides behind the other's Viterbi kernel.
 *
 * The two halves are also entry points of their own (what the GNU Radio blocks call): demod_submit() is the first half plus
 * the demod kernel, decode_submit() everything from the Viterbi kernel on, fed from host memory, device memory or an LLR
 * lease -- the LLR buffer of a collected demod batch, detached from its slot (llr_export) and entered in a process-wide table
 * (g_leases, one mutex) so that another context on the same GPU, the decode block's, reads it in place.  Retired leases hand
 * their buffers back to the producer's free list; nothing on that path calls cudaFree.
 */
#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <deque>
#include <mutex>
#include <thread>
#include <unordered_map>
#include <vector>
#include "common.cuh"
#include <nvtx3/nvToolsExt.h> /* header-only; ranges show up in Nsight Systems timelines, cost nothing without a tool attached */

extern "C" void w80211_fill_tables(W80211Tables* t);

namespace {

/* one NVTX range per host-side stage of a batch (tracing) */
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};
struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
};

enum { EV_START, EV_COPIED, EV_H2D, EV_HDR0, EV_HDR, EV_HDRD, EV_PLAN, EV_DEMOD, EV_VIT0, EV_VIT, EV_TB, EV_POST, EV_END, EV_COUNT };

struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[EV_COUNT] = {};
    int max_ncbps = 0; /* largest symbol of the batch, in LLRs: sizes the demod kernel's staging rows */
    bool surv_is_shared = false; /* this batch's survivors are in ctx->surv_shared (large batch), not in surv */
    bool busy = false;
    /* rx_submit returns once the input copies and the header kernel are enqueued; the rest of the batch (host plan, demod,
     * decode) is enqueued by a later API call, when the descriptors have arrived (run_continuation) */
    int phase = 0;    /* 0 idle, 1 header stage enqueued, 2 everything enqueued */
    int cont_err = 0; /* result of the continuation, reported by rx_collect */
    const float2* c_iq[2] = {nullptr, nullptr};
    const w80211_pkt_in* c_pk = nullptr;
    int c_chan_stride = 64;
    /* device */
    DevBuf iq[2], iq_raw[2], pkts, chan_in, desc, chan, symprefix, llr, surv, bits, bytes, mpdus, counters, slot_pkt, pkt_slot, grp,
        dec_list, psdu_off;
    /* pinned */
    PinBuf h_desc, h_plan, h_mpdus, h_counters;
    /* batch in flight */
    int npkt = 0, nant = 1, ngroups = 0, ndec = 0, nsymtot = 0;
    int cr_first[5] = {0, 0, 0, 0, 0};
    size_t plan_o_psdu = 0; /* offset of the psdu_off array inside h_plan */
    bool did_demod = false, did_decode = false, in_host = true;
    uint64_t n_llr = 0, n_bytes = 0, trellis_steps = 0;
    uint32_t mpdu_cap = 0;   /* records the device / pinned buffers of this batch can hold */
    uint32_t mpdu_eager = 0; /* records copied to the host together with the counters */
    uint8_t* post_bytes = nullptr; /* where the post kernel writes PSDU bytes (the caller's device buffer or s.bytes) */
    const float* llr_src = nullptr; /* LLRs the Viterbi kernels read: s.llr.p, the caller's device buffer, or a lease's */
    uint64_t lease = 0;             /* lease this batch reads (decode_submit with W80211_MEM_LEASE), released by collect */
    std::vector<uint64_t> v_src_off; /* decode_batch: the caller's own LLR offsets, handed back with the descriptors */
    uint32_t launches = 0;
    w80211_batch_out* out = nullptr;
    std::vector<uint32_t> h_psdu_off;
    /* host plan scratch, kept between batches (a fresh 0.5 MB vector per array and batch costs page faults) */
    std::vector<int> v_symprefix, v_dec, v_slot_pkt, v_pkt_slot;
    std::vector<uint32_t> v_key;
    std::vector<int4> v_grp;
    w80211_timing timing = {};
    float host_plan_ms = 0.f;
};

/* buffers of the signal stage (w80211_signal_batch) */
struct SigStage {
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {};
    DevBuf iq[2], syncs, res, chan_all, acc, out_iq[2], out_chan;
    PinBuf h_res, h_acc;
    /* detection stage */
    DevBuf det_ac, det_conj, det_sig, det_cnt, det_starts, det_ends, det_tmp, det_ev, det_res;
    DevBuf str_sig, str_ac, str_conj; /* w80211_detect_stream: samples and presiso streams of the buffer being examined */
    PinBuf h_edges, h_ev, h_dres;
    unsigned int edge_cap = 4096; /* run edges the lists are sized for: one and a half times what the last call found */
};

} // namespace

struct w80211_ctx {
    int device = 0;
    int nslots = 2;
    int vit_mode = W80211_CFG_VITERBI_AUTO; /* w80211_cfg.flags & W80211_CFG_VITERBI_MASK */
    int warp_max_pkts = 7168;               /* AUTO: a code rate with fewer packets than this in a batch takes the warp-per-packet kernels */
    cudaEvent_t origin = nullptr;    /* W80211_TRACE=1: time zero of the per-batch timelines printed by collect_slot */
    std::deque<int> pending;         /* slots in phase 1, in submission order */
    /* The input copies of all slots go through this one stream, so they reach the device one batch after the other at the
     * full rate of the link; on the slots' own streams the copy engine would interleave them and every batch in flight
     * would receive its last byte at about the same time. */
    cudaStream_t cstream = nullptr;
    cudaEvent_t last_tail = nullptr; /* EV_POST of the batch whose decode was enqueued last (see enqueue_front) */
    /* Survivors live from a batch's Viterbi kernel to its traceback, and large batches run those one after the other whatever
     * slot they are in (a Viterbi kernel fills the GPU), so from W80211_SURV_SHARE_BYTES up (default 1 GiB) all slots use this
     * one buffer, in turn: 29.8 GB once instead of once per slot at 113 664 C4 packets.  Smaller batches keep a buffer per slot
     * and may overlap. */
    DevBuf surv_shared;
    size_t surv_share_min = (size_t)1 << 30;
    cudaEvent_t last_shared_tail = nullptr; /* EV_POST of the last batch that used surv_shared */
    W80211Tables* d_tab = nullptr;
    /* LLR buffers this context has handed out as leases come back here when they retire (under g_lease_mu) */
    std::vector<DevBuf> llr_free;
    std::deque<uint64_t> leases; /* handles exported by this context, oldest first */
    size_t lease_bytes = (size_t)4 << 30; /* W80211_LEASE_BYTES: device memory the leases beyond the newest six may hold */
    Slot slots[4];
    SigStage sig;
    w80211_stats stats = {};
    std::string err;
};

namespace {

int fail(w80211_ctx* c, int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    return code;
}

#define CK(call)                                                                                              \
    do {                                                                                                      \
        cudaError_t e_ = (call);                                                                              \
        if (e_ != cudaSuccess) return fail(ctx, W80211_E_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
    } while (0)

int dev_reserve(w80211_ctx* ctx, DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return 0;
    if (b.p) CK(cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
    size_t want = bytes + bytes / 8 + 4096;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        want = bytes;
        e = cudaMalloc(&b.p, want);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(ctx, W80211_E_NOMEM, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
        }
    }
    b.cap = want;
    return 0;
}

int pin_reserve(w80211_ctx* ctx, PinBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return 0;
    if (b.p) CK(cudaFreeHost(b.p));
    b.p = nullptr;
    b.cap = 0;
    size_t want = bytes + bytes / 4 + 4096;
    CK(cudaHostAlloc(&b.p, want, cudaHostAllocDefault));
    b.cap = want;
    return 0;
}

#define RS(call)            \
    do {                    \
        int r_ = (call);    \
        if (r_ < 0) return r_; \
    } while (0)

/* ---- LLR leases: a demod batch's LLR buffer, detached from its slot so that another context on the same GPU can decode
 * from it (w80211rx.h).  One process-wide table; every access under g_lease_mu.  The buffer of a retired lease goes back
 * to its producer's free list (or is freed when the producer is gone), never to cudaFree on the hot path: cudaFree
 * synchronises the device. */
struct Lease {
    int device = 0;
    DevBuf buf;
    uint64_t nfloats = 0;
    w80211_ctx* owner = nullptr; /* nullptr: the producer has been destroyed */
    int readers = 0;             /* decode batches in flight that read it */
    bool retired = false;        /* no longer valid for new readers; freed when readers drops to 0 */
};
std::mutex g_lease_mu;
std::unordered_map<uint64_t, Lease> g_leases;
uint64_t g_lease_next = 1;
int g_lease_consumers = 0;

/* g_lease_mu held */
void lease_free_locked(uint64_t id)
{
    auto it = g_leases.find(id);
    if (it == g_leases.end()) return;
    Lease& L = it->second;
    if (L.owner)
        L.owner->llr_free.push_back(L.buf);
    else if (L.buf.p) {
        int cur = -1;
        cudaGetDevice(&cur);
        cudaSetDevice(L.device);
        cudaFree(L.buf.p);
        if (cur >= 0) cudaSetDevice(cur);
    }
    g_leases.erase(it);
}

/* the slot's LLR buffer, at least `bytes`: kept, or taken from the context's free list, or allocated */
int llr_reserve(w80211_ctx* ctx, Slot& s, size_t bytes)
{
    if (bytes <= s.llr.cap) return 0;
    {
        std::lock_guard<std::mutex> lk(g_lease_mu);
        int best = -1;
        for (int i = 0; i < (int)ctx->llr_free.size(); i++)
            if (ctx->llr_free[i].cap >= bytes && (best < 0 || ctx->llr_free[i].cap < ctx->llr_free[best].cap)) best = i;
        if (best >= 0) {
            if (s.llr.p) ctx->llr_free.push_back(s.llr);
            s.llr = ctx->llr_free[best];
            ctx->llr_free.erase(ctx->llr_free.begin() + best);
            return 0;
        }
    }
    return dev_reserve(ctx, s.llr, bytes);
}

void lease_release(uint64_t id)
{
    if (!id) return;
    std::lock_guard<std::mutex> lk(g_lease_mu);
    auto it = g_leases.find(id);
    if (it == g_leases.end()) return;
    if (--it->second.readers <= 0 && it->second.retired) lease_free_locked(id);
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

/* LLRs one packet needs for `trellis` steps at code rate cr (depuncturing of decode_impl.cc:210-229) */
inline long long llr_needed(int trellis, int cr)
{
    static const int per[4] = {1, 2, 3, 5}, cnt[4] = {2, 3, 4, 6};
    long long full = trellis / per[cr], rem = trellis % per[cr];
    long long n = full * cnt[cr];
    /* phase 0 takes 2, every later phase 1 */
    if (rem > 0) n += 2 + (rem - 1);
    return n;
}

/* Host plan after the header stage.  Fills llr offsets in h_desc, uploads the plan arrays. */
int plan_and_upload(w80211_ctx* ctx, Slot& s, w80211_desc* hd, bool assign_llr, bool do_decode)
{
    NvtxRange nvtx_("w80211 host plan + upload");
    const int npkt = s.npkt;
    /* One pass over the descriptors (12 MB for a full batch; the later loops only touch a 4-byte key per packet): LLR
     * layout, symbol prefix sums, and the packets to decode with their sort key (code rate ascending, trellis descending). */
    uint64_t off = 0;
    static const int spc2 = getenv("W80211_SPC2") ? std::max(1, atoi(getenv("W80211_SPC2"))) : W_DEMOD_SPC2;
    std::vector<int>& symprefix = s.v_symprefix;
    symprefix.assign((size_t)npkt + 1, 0);
    std::vector<int>& dec = s.v_dec;
    std::vector<uint32_t>& key = s.v_key;
    dec.clear();
    key.assign((size_t)npkt, 0);
    s.trellis_steps = 0;
    s.max_ncbps = 0;
    bool sorted = true;
    uint32_t last_key = 0;
    for (int p = 0; p < npkt; p++) {
        w80211_desc& d = hd[p];
        int nsym = 0;
        if (d.status == W80211_ST_OK && d.total > 0) {
            s.max_ncbps = std::max(s.max_ncbps, d.ncbps);
            if (assign_llr) {
                d.llr_off = (int64_t)off;
                off += align_up((size_t)d.total, 4) + 16;
            }
            nsym = d.nsym;
        } else if (assign_llr) {
            d.llr_off = -1;
        }
        /* demod work items: runs of at most spc symbols of one packet.  Two-stream packets take short runs, so that the 32
         * groups resident on an SM work on a few packets and share their channel matrices (4 KB each) in L1 */
        const int spc = (s.nant == 2 && d.nss == 2) ? spc2 : W_DEMOD_SPC;
        symprefix[p + 1] = symprefix[p] + (nsym + spc - 1) / spc;
        if (!do_decode || d.status != W80211_ST_OK) continue;
        /* decode_impl.cc:93: len > 4095 or trellis > 32782 -> skipped */
        if (d.len > W80211_B_MAX || d.len < 0 || d.trellis > W80211_T_MAX || d.trellis <= 0) continue;
        if (d.cr < 0 || d.cr > 3 || d.llr_off < 0 || (d.llr_off & 3)) continue;
        if (llr_needed(d.trellis, d.cr) > (long long)d.total) continue;
        const uint32_t k = ((uint32_t)d.cr << 16) | (uint32_t)(0xffff - d.trellis);
        key[p] = k;
        sorted = sorted && k >= last_key;
        last_key = k;
        dec.push_back(p);
        s.trellis_steps += (uint64_t)d.trellis;
    }
    if (assign_llr) s.n_llr = off;
    s.nsymtot = symprefix[npkt];
    if (!sorted) std::stable_sort(dec.begin(), dec.end(), [&](int a2, int b2) { return key[a2] < key[b2]; }); /* homogeneous batches: nothing to do */
    s.ndec = (int)dec.size();
    std::vector<int>&slot_pkt = s.v_slot_pkt, &pkt_slot = s.v_pkt_slot;
    std::vector<int4>& grp = s.v_grp;
    slot_pkt.clear();
    grp.clear();
    pkt_slot.assign((size_t)npkt, -1);
    s.h_psdu_off.assign(npkt, 0);
    uint64_t row = 0, wrow = 0, boff = 0;
    for (int c = 0; c <= 4; c++) s.cr_first[c] = 0;
    for (size_t i = 0; i < dec.size();) {
        const int cr = (int)(key[dec[i]] >> 16);
        const int maxT = 0xffff - (int)(key[dec[i]] & 0xffffu);
        int4 G;
        G.x = (int)row;
        G.y = (maxT + 1) & ~1;
        G.z = (int)wrow;
        G.w = cr;
        int k = 0;
        for (; k < 32 && i < dec.size() && (int)(key[dec[i]] >> 16) == cr; k++, i++) {
            const int p = dec[i];
            pkt_slot[p] = (int)slot_pkt.size();
            slot_pkt.push_back(p);
            s.h_psdu_off[p] = (uint32_t)boff; /* output bytes in decode order */
            const int T = 0xffff - (int)(key[p] & 0xffffu);
            boff += align_up((size_t)((T - 16 + 7) / 8) + 8, 16);
        }
        for (; k < 32; k++) slot_pkt.push_back(-1);
        for (int c = cr + 1; c <= 4; c++) s.cr_first[c] = (int)grp.size() + 1;
        grp.push_back(G);
        row += (uint64_t)G.y;
        wrow += (uint64_t)((maxT + 31) >> 5);
        if (row > 0x7fffffffull) return fail(ctx, W80211_E_CAPACITY, "batch too large: survivor rows overflow");
        if (boff > 0xffffffffull) return fail(ctx, W80211_E_CAPACITY, "batch too large: output bytes overflow 4 GiB");
    }
    s.n_bytes = boff;
    s.ngroups = (int)grp.size();

    /* device buffers */
    RS(dev_reserve(ctx, s.symprefix, (size_t)(npkt + 1) * 4));
    if (assign_llr) RS(llr_reserve(ctx, s, (size_t)(s.n_llr + 64) * 4));
    if (do_decode) {
        const size_t surv_bytes = (size_t)row * 32 * 8 + 1024;
        s.surv_is_shared = ctx->nslots > 1 && surv_bytes >= ctx->surv_share_min;
        /* growing the shared buffer frees the old one, which waits for every kernel still using it (cudaFree synchronises) */
        RS(dev_reserve(ctx, s.surv_is_shared ? ctx->surv_shared : s.surv, surv_bytes));
        RS(dev_reserve(ctx, s.bits, (size_t)wrow * 32 * 4 + 1024));
        RS(dev_reserve(ctx, s.bytes, (size_t)boff + 64));
        RS(dev_reserve(ctx, s.slot_pkt, slot_pkt.size() * 4 + 4));
        RS(dev_reserve(ctx, s.pkt_slot, (size_t)npkt * 4 + 4));
        RS(dev_reserve(ctx, s.grp, grp.size() * 16 + 16));
        RS(dev_reserve(ctx, s.dec_list, dec.size() * 4 + 4));
        RS(dev_reserve(ctx, s.psdu_off, (size_t)npkt * 4 + 4));
        RS(dev_reserve(ctx, s.counters, 64));
    }
    /* pinned staging for the plan, one block */
    size_t sz_sym = (size_t)(npkt + 1) * 4, sz_sp = slot_pkt.size() * 4, sz_ps = (size_t)npkt * 4, sz_g = grp.size() * 16,
           sz_d = dec.size() * 4, sz_po = (size_t)npkt * 4;
    size_t o_sym = 0, o_sp = align_up(o_sym + sz_sym, 16), o_ps = align_up(o_sp + sz_sp, 16), o_g = align_up(o_ps + sz_ps, 16),
           o_d = align_up(o_g + sz_g, 16), o_po = align_up(o_d + sz_d, 16), tot = align_up(o_po + sz_po, 16);
    RS(pin_reserve(ctx, s.h_plan, tot));
    char* hp = (char*)s.h_plan.p;
    memcpy(hp + o_sym, symprefix.data(), sz_sym);
    if (sz_sp) memcpy(hp + o_sp, slot_pkt.data(), sz_sp);
    memcpy(hp + o_ps, pkt_slot.data(), sz_ps);
    if (sz_g) memcpy(hp + o_g, grp.data(), sz_g);
    if (sz_d) memcpy(hp + o_d, dec.data(), sz_d);
    memcpy(hp + o_po, s.h_psdu_off.data(), sz_po);
    s.plan_o_psdu = o_po;
    cudaStream_t st = s.stream;
    /* host-fed batches: later batches' input copies are already queued on the copy engine, so the plan goes up through a
     * kernel that reads the pinned buffers; resident batches: the SMs may be busy, the copy engine is not */
    const bool by_kernel = s.in_host && s.did_demod;
    auto up = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
        if (bytes == 0) return cudaSuccess;
        if (by_kernel) {
            w_launch_upload(st, dst, src, bytes);
            s.launches += 1;
            return cudaGetLastError();
        }
        return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st);
    };
    CK(up(s.symprefix.p, hp + o_sym, sz_sym));
    if (do_decode) {
        CK(up(s.slot_pkt.p, hp + o_sp, sz_sp));
        CK(up(s.pkt_slot.p, hp + o_ps, sz_ps));
        CK(up(s.grp.p, hp + o_g, sz_g));
        CK(up(s.dec_list.p, hp + o_d, sz_d));
        CK(up(s.psdu_off.p, hp + o_po, sz_po));
        CK(cudaMemsetAsync(s.counters.p, 0, 64, st));
    }
    /* descriptors with their LLR offsets back to the device */
    CK(up(s.desc.p, hd, (size_t)npkt * sizeof(w80211_desc)));
    return 0;
}

int enqueue_decode_tail(w80211_ctx* ctx, Slot& s)
{
    NvtxRange nvtx_("w80211 enqueue viterbi/traceback/post");
    cudaStream_t st = s.stream;
    w80211_batch_out* out = s.out;
    uint2* surv = (uint2*)(s.surv_is_shared ? ctx->surv_shared.p : s.surv.p);
    /* the shared survivor buffer is free once the traceback of the batch that used it last has run (a batch that came in
     * through rx_submit has waited for that already, before its demod kernel; a decode-only batch has not) */
    if (s.surv_is_shared && ctx->last_shared_tail && ctx->last_shared_tail != s.ev[EV_TB]) CK(cudaStreamWaitEvent(st, ctx->last_shared_tail, 0));
    CK(cudaEventRecord(s.ev[EV_VIT0], st));
    int warp_mask = 0;
    for (int cr = 0; cr < 4; cr++) {
        const int npk = (s.cr_first[cr + 1] - s.cr_first[cr]) * 32;
        if (ctx->vit_mode == W80211_CFG_VITERBI_WARP || (ctx->vit_mode == W80211_CFG_VITERBI_AUTO && npk < ctx->warp_max_pkts)) warp_mask |= 1 << cr;
    }
    int nvit = w_launch_viterbi(st, s.cr_first, (const int4*)s.grp.p, (const int*)s.slot_pkt.p, (const w80211_desc*)s.desc.p,
                                s.llr_src ? s.llr_src : (const float*)s.llr.p, surv, warp_mask);
    CK(cudaGetLastError());
    CK(cudaEventRecord(s.ev[EV_VIT], st));
    nvit += w_launch_traceback(st, s.cr_first, (const int4*)s.grp.p, (const int*)s.slot_pkt.p, (const w80211_desc*)s.desc.p,
                               surv, (uint32_t*)s.bits.p, warp_mask);
    CK(cudaGetLastError());
    CK(cudaEventRecord(s.ev[EV_TB], st));
    if (s.surv_is_shared) ctx->last_shared_tail = s.ev[EV_TB];
    /* MPDU records: one per packet is the rule, a VHT A-MPDU has one per delimiter walked (up to trellis / 32 + 1).  The
     * buffers start at 1.5 per packet; a batch that produces more is finished by collect_slot(), which knows the exact
     * count by then, grows the buffers and runs the post kernel once more (launch_post_again). */
    uint32_t internal_cap = (uint32_t)std::min<size_t>((size_t)s.ndec + (size_t)s.ndec / 2 + 64, 0xffffffffu);
    RS(dev_reserve(ctx, s.mpdus, (size_t)internal_cap * sizeof(w80211_mpdu)));
    RS(pin_reserve(ctx, s.h_mpdus, (size_t)internal_cap * sizeof(w80211_mpdu)));
    RS(pin_reserve(ctx, s.h_counters, 64));
    s.mpdu_cap = internal_cap;
    /* Results for a caller with device buffers are written in place (PSDU bytes by the post kernel) or uploaded from the
     * pinned host copies (descriptors, offsets): no device-to-device copy kernel, which would have to wait for SMs while
     * another slot's Viterbi kernel holds all of them and so delay rx_collect by half that kernel's duration. */
    const bool bytes_in_place = out->bytes && out->mem == W80211_MEM_DEVICE && ((uintptr_t)out->bytes & 15u) == 0;
    if (out->bytes && out->bytes_capacity < s.n_bytes)
        return fail(ctx, W80211_E_CAPACITY, "bytes_capacity %llu < %llu needed", (unsigned long long)out->bytes_capacity,
                    (unsigned long long)s.n_bytes);
    s.post_bytes = bytes_in_place ? out->bytes : (uint8_t*)s.bytes.p;
    w_launch_post(st, ctx->d_tab, s.ndec, (const int4*)s.grp.p, (const int*)s.pkt_slot.p, (const int*)s.dec_list.p,
                  (const w80211_desc*)s.desc.p, (const uint32_t*)s.bits.p, (const uint32_t*)s.psdu_off.p, s.post_bytes,
                  (w80211_mpdu*)s.mpdus.p, internal_cap, (unsigned int*)s.counters.p);
    CK(cudaGetLastError());
    CK(cudaEventRecord(s.ev[EV_POST], st));
    ctx->last_tail = s.ev[EV_POST];
    if (s.ngroups > 0) s.launches += 1 + nvit;
    /* results: the counters and one record per packet now, the rest (A-MPDUs with several subframes) when collect_slot()
     * has read the count */
    s.mpdu_eager = (uint32_t)std::min<size_t>((size_t)s.ndec + 64, internal_cap);
    CK(cudaMemcpyAsync(s.h_counters.p, s.counters.p, 64, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(s.h_mpdus.p, s.mpdus.p, (size_t)s.mpdu_eager * sizeof(w80211_mpdu), cudaMemcpyDeviceToHost, st));
    if (out->bytes && !bytes_in_place && s.n_bytes)
        CK(cudaMemcpyAsync(out->bytes, s.bytes.p, s.n_bytes,
                           out->mem == W80211_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
    if (out->psdu_off) {
        if (out->mem == W80211_MEM_DEVICE)
            CK(cudaMemcpyAsync(out->psdu_off, (const char*)s.h_plan.p + s.plan_o_psdu, (size_t)s.npkt * 4, cudaMemcpyHostToDevice, st));
        else
            memcpy(out->psdu_off, s.h_psdu_off.data(), (size_t)s.npkt * 4);
    }
    return 0;
}

int finish_outputs(w80211_ctx* ctx, Slot& s)
{
    cudaStream_t st = s.stream;
    w80211_batch_out* out = s.out;
    const w80211_desc* hd = (const w80211_desc*)s.h_desc.p;
    if (out->desc) {
        if (out->mem == W80211_MEM_DEVICE)
            CK(cudaMemcpyAsync(out->desc, hd, (size_t)s.npkt * sizeof(w80211_desc), cudaMemcpyHostToDevice, st));
        else
            memcpy(out->desc, hd, (size_t)s.npkt * sizeof(w80211_desc));
    }
    if (out->want_llr && out->llr) {
        if (out->llr_capacity < s.n_llr) return fail(ctx, W80211_E_CAPACITY, "llr_capacity %llu < %llu needed",
                                                     (unsigned long long)out->llr_capacity, (unsigned long long)s.n_llr);
        if (s.n_llr)
            CK(cudaMemcpyAsync(out->llr, s.llr.p, s.n_llr * 4,
                               out->mem == W80211_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
    }
    CK(cudaEventRecord(s.ev[EV_END], st));
    return 0;
}

/* front half shared by demod_batch and rx_submit */
/* first half of a batch: input copies (or nothing for device input), integer IQ expansion, header kernel, descriptors
 * back to the host.  Does not wait. */
int enqueue_input_header(w80211_ctx* ctx, Slot& s, const w80211_batch_in* in, w80211_batch_out* out, bool do_decode)
{
    NvtxRange nvtx_("w80211 enqueue input copies + header");
    if (!in || !out || in->npkt < 0 || (in->nant != 1 && in->nant != 2) || !in->iq[0] || (in->nant == 2 && !in->iq[1]) ||
        (in->npkt > 0 && (!in->pkts || !in->chan)) || (in->iq_format != W80211_IQ_FC32 && in->iq_format != W80211_IQ_SC16 && in->iq_format != W80211_IQ_SC8))
        return fail(ctx, W80211_E_ARG, "bad batch_in");
    const bool sc16 = in->iq_format != W80211_IQ_FC32; /* integer wire format, expanded on the GPU */
    const int raw_bytes = in->iq_format == W80211_IQ_SC8 ? 2 : 4;
    if (s.busy) return fail(ctx, W80211_E_STATE, "slot busy: collect it first");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = s.stream;
    const int npkt = in->npkt;
    s.npkt = npkt;
    s.nant = in->nant;
    s.out = out;
    s.did_demod = true;
    s.did_decode = do_decode;
    s.launches = 0;
    s.llr_src = nullptr;
    s.lease = 0;
    s.in_host = in->mem == W80211_MEM_HOST;
    out->n_mpdu = out->n_crc_ok = 0;
    out->n_bytes = out->n_llr = out->psdu_bits_ok = 0;
    const int chan_stride = in->nant == 2 ? 520 : 64;
    CK(cudaEventRecord(s.ev[EV_START], st));
    const float2* d_iq[2] = {nullptr, nullptr};
    const w80211_pkt_in* d_pk;
    const float2* d_chin;
    if (s.in_host) {
        cudaStream_t cs = ctx->cstream ? ctx->cstream : st;
        RS(dev_reserve(ctx, s.pkts, (size_t)npkt * sizeof(w80211_pkt_in) + 16));
        RS(dev_reserve(ctx, s.chan_in, (size_t)npkt * 512 + 16));
        for (int a = 0; a < in->nant; a++) {
            RS(dev_reserve(ctx, s.iq[a], (size_t)in->nsamples * 8 + 1024));
            if (sc16) RS(dev_reserve(ctx, s.iq_raw[a], (size_t)in->nsamples * raw_bytes + 1024));
        }
        if (npkt) {
            CK(cudaMemcpyAsync(s.pkts.p, in->pkts, (size_t)npkt * sizeof(w80211_pkt_in), cudaMemcpyHostToDevice, cs));
            CK(cudaMemcpyAsync(s.chan_in.p, in->chan, (size_t)npkt * 512, cudaMemcpyHostToDevice, cs));
        }
        d_pk = (const w80211_pkt_in*)s.pkts.p;
        d_chin = (const float2*)s.chan_in.p;
        for (int a = 0; a < in->nant; a++) {
            if (sc16)
                CK(cudaMemcpyAsync(s.iq_raw[a].p, in->iq[a], (size_t)in->nsamples * raw_bytes, cudaMemcpyHostToDevice, cs));
            else
                CK(cudaMemcpyAsync(s.iq[a].p, in->iq[a], (size_t)in->nsamples * 8, cudaMemcpyHostToDevice, cs));
            d_iq[a] = (const float2*)s.iq[a].p;
        }
        if (cs != st) {
            CK(cudaEventRecord(s.ev[EV_COPIED], cs));
            CK(cudaStreamWaitEvent(st, s.ev[EV_COPIED], 0));
        }
        if (sc16) {
            w_launch_sc16_to_fc32(st, npkt, in->nant, d_pk, s.iq_raw[0].p, s.iq_raw[1].p, (float2*)s.iq[0].p, (float2*)s.iq[1].p,
                                  in->nsamples, raw_bytes);
            CK(cudaGetLastError());
            if (npkt) s.launches += 1;
        }
    } else {
        d_pk = in->pkts;
        d_chin = (const float2*)in->chan;
        for (int a = 0; a < in->nant; a++) {
            if (sc16) {
                RS(dev_reserve(ctx, s.iq[a], (size_t)in->nsamples * 8 + 1024));
                d_iq[a] = (const float2*)s.iq[a].p;
            } else {
                d_iq[a] = (const float2*)in->iq[a];
            }
        }
        if (sc16) {
            w_launch_sc16_to_fc32(st, npkt, in->nant, d_pk, in->iq[0], in->iq[1], (float2*)s.iq[0].p, (float2*)s.iq[1].p, in->nsamples,
                                  raw_bytes);
            CK(cudaGetLastError());
            if (npkt) s.launches += 1;
        }
    }
    CK(cudaEventRecord(s.ev[EV_H2D], st));
    RS(dev_reserve(ctx, s.desc, (size_t)npkt * sizeof(w80211_desc) + 16));
    RS(dev_reserve(ctx, s.chan, (size_t)npkt * chan_stride * 8 + 16));
    RS(pin_reserve(ctx, s.h_desc, (size_t)npkt * sizeof(w80211_desc) + 16));
    CK(cudaEventRecord(s.ev[EV_HDR0], st));
    w_launch_header(st, ctx->d_tab, npkt, in->nant, d_iq[0], d_iq[1], in->nsamples, d_pk, d_chin, (w80211_desc*)s.desc.p,
                    (float2*)s.chan.p, chan_stride, in->mu_pos ? 1 : 0);
    CK(cudaGetLastError());
    if (npkt) s.launches += 1;
    CK(cudaEventRecord(s.ev[EV_HDR], st));
    if (npkt) CK(cudaMemcpyAsync(s.h_desc.p, s.desc.p, (size_t)npkt * sizeof(w80211_desc), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(s.ev[EV_HDRD], st));
    s.c_iq[0] = d_iq[0];
    s.c_iq[1] = d_iq[1];
    s.c_pk = d_pk;
    s.c_chan_stride = chan_stride;
    return 0;
}

/* second half of the front: the descriptors are on the host (EV_HDRD has completed); plan, then demod */
int continue_front(w80211_ctx* ctx, Slot& s)
{
    NvtxRange nvtx_("w80211 enqueue demod");
    cudaStream_t st = s.stream;
    const int npkt = s.npkt;
    const bool do_decode = s.did_decode;
    const float2* const* d_iq = s.c_iq;
    const w80211_pkt_in* d_pk = s.c_pk;
    const int chan_stride = s.c_chan_stride;
    w80211_desc* hd = (w80211_desc*)s.h_desc.p;
    const auto tp0 = std::chrono::steady_clock::now();
    RS(plan_and_upload(ctx, s, hd, true, do_decode));
    s.host_plan_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - tp0).count();
    /* Several batches in flight: this batch's demod kernel (and with it its Viterbi kernel) waits for the traceback and
     * post kernels of the batch enqueued before it, i.e. for that batch's whole decode.  The GPU then runs
     *     Viterbi A | traceback A, post A | demod B | Viterbi B | traceback B, post B | demod A' | ...
     * one big kernel at a time -- sharing the SMs with a Viterbi kernel slows both sides down by more than running
     * them in turn (it uses 84 % of the issue slots; profiles/experiments/r1_overlap_shared_viterbi_stream.txt) -- and
     * the slots alternate instead of falling into lockstep (both Viterbi kernels interleaved, both tails at once, then
     * an idle GPU while the host collects and plans): the header kernel, the host plan and rx_collect of one slot all
     * happen while the other slot's Viterbi kernel runs. */
    if (ctx->last_tail && ctx->last_tail != s.ev[EV_POST]) CK(cudaStreamWaitEvent(st, ctx->last_tail, 0));
    CK(cudaEventRecord(s.ev[EV_PLAN], st));
    w_launch_demod(st, ctx->d_tab, s.nsymtot, npkt, s.nant, d_iq[0], d_iq[1], d_pk, (const w80211_desc*)s.desc.p,
                   (const int*)s.symprefix.p, (const float2*)s.chan.p, chan_stride, (float*)s.llr.p, s.max_ncbps);
    CK(cudaGetLastError());
    if (s.nsymtot) s.launches += 1;
    CK(cudaEventRecord(s.ev[EV_DEMOD], st));
    return 0;
}

/* The post kernel counted `need` MPDU records but could store only s.mpdu_cap of them: grow the buffers to the exact
 * count and run it again (its inputs are still in the slot's buffers; it rewrites the same bytes). */
int launch_post_again(w80211_ctx* ctx, Slot& s, uint32_t need)
{
    cudaStream_t st = s.stream;
    RS(dev_reserve(ctx, s.mpdus, (size_t)need * sizeof(w80211_mpdu)));
    RS(pin_reserve(ctx, s.h_mpdus, (size_t)need * sizeof(w80211_mpdu)));
    s.mpdu_cap = need;
    CK(cudaMemsetAsync(s.counters.p, 0, 64, st));
    w_launch_post(st, ctx->d_tab, s.ndec, (const int4*)s.grp.p, (const int*)s.pkt_slot.p, (const int*)s.dec_list.p,
                  (const w80211_desc*)s.desc.p, (const uint32_t*)s.bits.p, (const uint32_t*)s.psdu_off.p, s.post_bytes,
                  (w80211_mpdu*)s.mpdus.p, need, (unsigned int*)s.counters.p);
    CK(cudaGetLastError());
    s.launches += 1;
    CK(cudaMemcpyAsync(s.h_counters.p, s.counters.p, 64, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(s.h_mpdus.p, s.mpdus.p, (size_t)need * sizeof(w80211_mpdu), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (((const unsigned int*)s.h_counters.p)[0] != need) return fail(ctx, W80211_E_STATE, "MPDU record count changed between two runs of the post kernel");
    return 0;
}

int collect_slot(w80211_ctx* ctx, Slot& s)
{
    NvtxRange nvtx_("w80211 collect");
    if (!s.busy) return fail(ctx, W80211_E_STATE, "slot idle: nothing to collect");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(s.stream));
    const auto tc0 = std::chrono::steady_clock::now();
    s.busy = false;
    if (s.lease) {
        lease_release(s.lease);
        s.lease = 0;
    }
    w80211_batch_out* out = s.out;
    const w80211_desc* hd = (const w80211_desc*)s.h_desc.p;
    out->n_llr = s.n_llr;
    w80211_stats& S = ctx->stats;
    S.packets_in += (uint64_t)s.npkt;
    S.kernel_launches += s.launches;
    if (s.did_demod) {
        for (int p = 0; p < s.npkt; p++) {
            switch (hd[p].status) {
            case W80211_ST_DROP_GATE: S.dropped_gate++; break;
            case W80211_ST_NDP: S.ndp++; break;
            case W80211_ST_UNSUPPORTED: S.unsupported++; break;
            case W80211_ST_BAD_INPUT: S.bad_input++; break;
            default: break;
            }
        }
    }
    if (s.did_decode) {
        const unsigned int* c = (const unsigned int*)s.h_counters.p;
        uint32_t n = c[0];
        if (n > s.mpdu_cap) RS(launch_post_again(ctx, s, n)); /* more subframes than the buffers held: now the count is known */
        else if (n > s.mpdu_eager) {
            CK(cudaMemcpyAsync((w80211_mpdu*)s.h_mpdus.p + s.mpdu_eager, (const w80211_mpdu*)s.mpdus.p + s.mpdu_eager,
                               (size_t)(n - s.mpdu_eager) * sizeof(w80211_mpdu), cudaMemcpyDeviceToHost, s.stream));
            CK(cudaStreamSynchronize(s.stream));
        }
        c = (const unsigned int*)s.h_counters.p;
        n = c[0];
        w80211_mpdu* hm = (w80211_mpdu*)s.h_mpdus.p;
        /* records arrive in completion order; put them in (pkt, off) order with one counting pass over the packets */
        auto rec_before = [](const w80211_mpdu& a, const w80211_mpdu& b) { return a.pkt != b.pkt ? a.pkt < b.pkt : a.off < b.off; };
        if (!std::is_sorted(hm, hm + n, rec_before)) {
            std::vector<uint32_t> first((size_t)s.npkt + 1, 0);
            bool in_range = true;
            for (uint32_t i = 0; i < n; i++) {
                if (hm[i].pkt >= (uint32_t)s.npkt) { in_range = false; break; }
                first[hm[i].pkt + 1]++;
            }
            if (in_range) {
                for (int p = 0; p < s.npkt; p++) first[(size_t)p + 1] += first[(size_t)p];
                std::vector<w80211_mpdu> tmp(hm, hm + n);
                std::vector<uint32_t> pos(first.begin(), first.end() - 1);
                for (uint32_t i = 0; i < n; i++) hm[pos[tmp[i].pkt]++] = tmp[i];
                for (int p = 0; p < s.npkt; p++)
                    if (first[(size_t)p + 1] - first[(size_t)p] > 1) std::sort(hm + first[(size_t)p], hm + first[(size_t)p + 1], rec_before);
            } else {
                std::sort(hm, hm + n, rec_before);
            }
        }
        out->n_mpdu = n;
        out->n_crc_ok = c[1];
        out->psdu_bits_ok = (uint64_t)c[2] | ((uint64_t)c[3] << 32);
        out->n_bytes = s.n_bytes;
        S.packets_decoded += (uint64_t)s.ndec;
        S.trellis_steps += s.trellis_steps;
        S.mpdu_total += n;
        S.mpdu_crc_ok += c[1];
        for (uint32_t i = 0; i < n; i++) {
            if (!hm[i].crc_ok) continue;
            const w80211_desc& d = hd[hm[i].pkt];
            if (d.format == W80211_F_L && d.mcs >= 0 && d.mcs < 8) S.ok_legacy[d.mcs]++;
            if (d.format == W80211_F_HT && d.mcs >= 0) S.ok_ht[d.mcs % 8]++;
            if (d.format == W80211_F_VHT && d.mcs >= 0 && d.mcs < 10) S.ok_vht[d.mcs]++;
        }
        if (out->mpdus) {
            if (out->mpdu_capacity < n) return fail(ctx, W80211_E_CAPACITY, "mpdu_capacity %u < %u", out->mpdu_capacity, n);
            if (out->mem == W80211_MEM_DEVICE) {
                CK(cudaMemcpy(out->mpdus, hm, (size_t)n * sizeof(w80211_mpdu), cudaMemcpyHostToDevice));
            } else {
                memcpy(out->mpdus, hm, (size_t)n * sizeof(w80211_mpdu));
            }
        }
    }
    /* stage timing */
    auto ms = [&](int a, int b) {
        float v = 0.f;
        cudaEventElapsedTime(&v, s.ev[a], s.ev[b]);
        return v;
    };
    w80211_timing& t = s.timing;
    memset(&t, 0, sizeof(t));
    if (s.did_demod) {
        t.h2d = ms(EV_START, EV_H2D);
        t.header = ms(EV_HDR0, EV_HDR);
        t.demod = ms(EV_PLAN, EV_DEMOD);
    }
    if (s.did_decode) {
        t.viterbi = ms(EV_VIT0, EV_VIT);
        t.traceback = ms(EV_VIT, EV_TB);
        t.post = ms(EV_TB, EV_POST);
        t.d2h = ms(EV_POST, EV_END);
    }
    t.total = ms(EV_START, EV_END);
    if (ctx->origin && s.did_demod && s.did_decode) {
        /* diagnostics: when each stage of this batch ended on the device, in ms since the context was created */
        static const int evs[] = {EV_START, EV_HDR0, EV_HDR, EV_PLAN, EV_DEMOD, EV_VIT0, EV_VIT, EV_TB, EV_POST, EV_END};
        static const char* names[] = {"start", "hdr0", "hdr", "plan", "demod", "vit0", "vit", "tb", "post", "end"};
        fprintf(stderr, "w80211 timeline slot %d:", (int)(&s - ctx->slots));
        for (int i = 0; i < 10; i++) {
            float v = 0.f;
            cudaEventElapsedTime(&v, ctx->origin, s.ev[evs[i]]);
            fprintf(stderr, " %s %.2f", names[i], v);
        }
        fprintf(stderr, "\n");
    }
    t.launches = s.launches;
    t.host_plan = s.host_plan_ms;
    t.host_collect = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - tc0).count();
    cudaGetLastError();
    return 0;
}


/* second half of a batch whose header stage has completed */
void run_continuation(w80211_ctx* ctx, Slot& s)
{
    int r = continue_front(ctx, s);
    if (r == 0 && s.did_decode) r = enqueue_decode_tail(ctx, s);
    if (r == 0) r = finish_outputs(ctx, s);
    s.cont_err = r;
    s.phase = 2;
}

/* Enqueue the second half of every pending batch, in submission order, whose descriptors have reached the host.
 * upto >= 0: wait for the descriptors of the batches up to and including slot `upto`; -2: of all of them. */
void progress(w80211_ctx* ctx, int upto)
{
    if (ctx->pending.empty()) return;
    cudaSetDevice(ctx->device); /* the caller may have used a context on another device since (one process, several GPUs) */
    while (!ctx->pending.empty()) {
        const int idx = ctx->pending.front();
        Slot& s = ctx->slots[idx];
        if (upto == -1) {
            if (cudaEventQuery(s.ev[EV_HDRD]) == cudaErrorNotReady) break;
        } else {
            cudaEventSynchronize(s.ev[EV_HDRD]);
        }
        ctx->pending.pop_front();
        run_continuation(ctx, s);
        if (idx == upto) upto = -1;
    }
    cudaGetLastError();
}

} // namespace

extern "C" {

int w80211_abi_version(void) { return W80211_ABI_VERSION; }

size_t w80211_sizeof(int what)
{
    switch (what) {
    case 0: return sizeof(w80211_pkt_in);
    case 1: return sizeof(w80211_batch_in);
    case 2: return sizeof(w80211_desc);
    case 3: return sizeof(w80211_mpdu);
    case 4: return sizeof(w80211_batch_out);
    case 5: return sizeof(w80211_stats);
    case 6: return sizeof(w80211_timing);
    case 7: return sizeof(w80211_cfg);
    case 8: return sizeof(w80211_sync);
    case 9: return sizeof(w80211_signal_in);
    case 10: return sizeof(w80211_signal_out);
    case 11: return sizeof(w80211_detect_state);
    case 12: return sizeof(w80211_detect_in);
    case 13: return sizeof(w80211_detect_out);
    case 14: return sizeof(w80211_decode_in);
    default: return 0;
    }
}

int w80211_create(const w80211_cfg* cfg, w80211_ctx** outp)
{
    if (!outp) return W80211_E_ARG;
    *outp = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        return W80211_E_CUDA; /* no CPU fallback */
    }
    w80211_ctx* ctx = new (std::nothrow) w80211_ctx();
    if (!ctx) return W80211_E_NOMEM;
    ctx->device = cfg ? cfg->device : 0;
    ctx->nslots = cfg && cfg->nslots > 0 ? std::min(cfg->nslots, 4) : 2;
    ctx->vit_mode = cfg ? (cfg->flags & W80211_CFG_VITERBI_MASK) : W80211_CFG_VITERBI_AUTO;
    if (const char* e = getenv("W80211_WARP_MAX_PKTS")) ctx->warp_max_pkts = atoi(e);
    if (const char* e = getenv("W80211_LEASE_BYTES")) ctx->lease_bytes = (size_t)strtoull(e, nullptr, 10);
    if (const char* e = getenv("W80211_SURV_SHARE_BYTES")) ctx->surv_share_min = (size_t)strtoull(e, nullptr, 10); /* 0: always shared */
    if (ctx->device < 0 || ctx->device >= ndev || cudaSetDevice(ctx->device) != cudaSuccess) {
        delete ctx;
        return W80211_E_ARG;
    }
    W80211Tables* ht = new W80211Tables;
    w80211_fill_tables(ht);
    bool ok = cudaMalloc(&ctx->d_tab, sizeof(W80211Tables)) == cudaSuccess &&
              cudaMemcpy(ctx->d_tab, ht, sizeof(W80211Tables), cudaMemcpyHostToDevice) == cudaSuccess;
    delete ht;
    ok = ok && w_demod_init_device() == 0;
    ok = ok && cudaStreamCreateWithFlags(&ctx->cstream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; ok && i < ctx->nslots; i++) {
        ok = cudaStreamCreateWithFlags(&ctx->slots[i].stream, cudaStreamNonBlocking) == cudaSuccess;
        for (int e = 0; ok && e < EV_COUNT; e++) ok = cudaEventCreate(&ctx->slots[i].ev[e]) == cudaSuccess;
    }
    if (ok && getenv("W80211_TRACE")) ok = cudaEventCreate(&ctx->origin) == cudaSuccess && cudaEventRecord(ctx->origin, 0) == cudaSuccess;
    if (!ok) {
        cudaGetLastError();
        w80211_destroy(ctx);
        return W80211_E_CUDA;
    }
    *outp = ctx;
    return W80211_OK;
}

void w80211_destroy(w80211_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->cstream) cudaStreamSynchronize(ctx->cstream);
    for (int i = 0; i < 4; i++) {
        Slot& s = ctx->slots[i];
        if (s.stream) cudaStreamSynchronize(s.stream);
        DevBuf* db[] = {&s.iq[0], &s.iq[1], &s.iq_raw[0], &s.iq_raw[1], &s.pkts, &s.chan_in, &s.desc, &s.chan, &s.symprefix, &s.llr, &s.surv, &s.bits,
                        &s.bytes, &s.mpdus, &s.counters, &s.slot_pkt, &s.pkt_slot, &s.grp, &s.dec_list, &s.psdu_off};
        for (DevBuf* b : db)
            if (b->p) cudaFree(b->p);
        PinBuf* pb[] = {&s.h_desc, &s.h_plan, &s.h_mpdus, &s.h_counters};
        for (PinBuf* b : pb)
            if (b->p) cudaFreeHost(b->p);
        for (int e = 0; e < EV_COUNT; e++)
            if (s.ev[e]) cudaEventDestroy(s.ev[e]);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    {
        SigStage& g = ctx->sig;
        DevBuf* db[] = {&g.iq[0], &g.iq[1], &g.syncs, &g.res, &g.chan_all, &g.acc, &g.out_iq[0], &g.out_iq[1], &g.out_chan,
                        &g.det_ac, &g.det_conj, &g.det_sig, &g.det_cnt, &g.det_starts, &g.det_ends, &g.det_tmp, &g.str_sig, &g.str_ac, &g.str_conj, &g.det_ev, &g.det_res};
        for (DevBuf* b : db)
            if (b->p) cudaFree(b->p);
        if (g.h_res.p) cudaFreeHost(g.h_res.p);
        if (g.h_acc.p) cudaFreeHost(g.h_acc.p);
        if (g.h_edges.p) cudaFreeHost(g.h_edges.p);
        if (g.h_ev.p) cudaFreeHost(g.h_ev.p);
        if (g.h_dres.p) cudaFreeHost(g.h_dres.p);
        for (auto& e : g.ev)
            if (e) cudaEventDestroy(e);
        if (g.stream) cudaStreamDestroy(g.stream);
    }
    {
        /* leases this context produced: those nobody reads are freed now, the others when their reader collects */
        std::lock_guard<std::mutex> lk(g_lease_mu);
        for (uint64_t id : ctx->leases) {
            auto it = g_leases.find(id);
            if (it == g_leases.end()) continue;
            it->second.owner = nullptr;
            it->second.retired = true;
            if (it->second.readers <= 0) lease_free_locked(id);
        }
        ctx->leases.clear();
        for (DevBuf& b : ctx->llr_free)
            if (b.p) cudaFree(b.p);
        ctx->llr_free.clear();
    }
    for (int i = 0; i < 4; i++) lease_release(ctx->slots[i].lease);
    if (ctx->cstream) cudaStreamDestroy(ctx->cstream);
    if (ctx->origin) cudaEventDestroy(ctx->origin);
    if (ctx->d_tab) cudaFree(ctx->d_tab);
    if (ctx->surv_shared.p) cudaFree(ctx->surv_shared.p);
    delete ctx;
}

const char* w80211_last_error(const w80211_ctx* ctx) { return ctx ? ctx->err.c_str() : "no context (CUDA device missing?)"; }

int w80211_stats_get(const w80211_ctx* ctx, w80211_stats* out)
{
    if (!ctx || !out) return W80211_E_ARG;
    *out = ctx->stats;
    return 0;
}

int w80211_timing_get(const w80211_ctx* ctx, int slot, w80211_timing* out)
{
    if (!ctx || !out || slot < 0 || slot >= ctx->nslots) return W80211_E_ARG;
    *out = ctx->slots[slot].timing;
    return 0;
}

void* w80211_stream(const w80211_ctx* ctx, int slot)
{
    if (!ctx || slot < 0 || slot >= ctx->nslots) return nullptr;
    return (void*)ctx->slots[slot].stream;
}

void* w80211_host_alloc(size_t bytes)
{
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}

void w80211_host_free(void* p)
{
    if (p) cudaFreeHost(p);
}

int w80211_rx_submit(w80211_ctx* ctx, int slot, const w80211_batch_in* in, w80211_batch_out* out)
{
    if (!ctx) return W80211_E_ARG;
    if (slot < 0 || slot >= ctx->nslots) return fail(ctx, W80211_E_ARG, "bad slot %d", slot);
    Slot& s = ctx->slots[slot];
    progress(ctx, -1);
    int r = enqueue_input_header(ctx, s, in, out, true);
    if (r < 0) {
        if (s.stream) cudaStreamSynchronize(s.stream);
        return r;
    }
    s.busy = true;
    s.phase = 1;
    s.cont_err = 0;
    ctx->pending.push_back(slot);
    return W80211_OK;
}

int w80211_rx_collect(w80211_ctx* ctx, int slot)
{
    if (!ctx) return W80211_E_ARG;
    if (slot < 0 || slot >= ctx->nslots) return fail(ctx, W80211_E_ARG, "bad slot %d", slot);
    Slot& s = ctx->slots[slot];
    if (!s.busy) return fail(ctx, W80211_E_STATE, "slot idle: nothing to collect");
    CK(cudaSetDevice(ctx->device));
    if (s.phase == 1) progress(ctx, slot);
    /* wait for this batch, and meanwhile enqueue the second half of any other batch whose descriptors arrive */
    while (!ctx->pending.empty()) {
        progress(ctx, -1);
        cudaError_t q = cudaStreamQuery(s.stream);
        if (q != cudaErrorNotReady) break;
        std::this_thread::sleep_for(std::chrono::microseconds(20));
    }
    if (s.cont_err < 0) {
        cudaStreamSynchronize(s.stream);
        s.busy = false;
        s.phase = 0;
        lease_release(s.lease);
        s.lease = 0;
        return s.cont_err;
    }
    int r = collect_slot(ctx, s);
    s.phase = 0;
    return r;
}

int w80211_rx_batch(w80211_ctx* ctx, const w80211_batch_in* in, w80211_batch_out* out)
{
    int r = w80211_rx_submit(ctx, 0, in, out);
    if (r < 0) return r;
    return w80211_rx_collect(ctx, 0);
}

int w80211_rx_poll(w80211_ctx* ctx, int slot)
{
    if (!ctx) return W80211_E_ARG;
    if (slot < 0 || slot >= ctx->nslots) return fail(ctx, W80211_E_ARG, "bad slot %d", slot);
    Slot& s = ctx->slots[slot];
    if (!s.busy) return fail(ctx, W80211_E_STATE, "slot idle: nothing to poll");
    CK(cudaSetDevice(ctx->device));
    progress(ctx, -1);
    if (s.phase != 2) return 0;
    if (s.cont_err < 0) return 1; /* collect will report it */
    const cudaError_t q = cudaStreamQuery(s.stream);
    if (q == cudaErrorNotReady) return 0;
    if (q != cudaSuccess) return fail(ctx, W80211_E_CUDA, "slot %d: %s", slot, cudaGetErrorString(q));
    return 1;
}

int w80211_demod_submit(w80211_ctx* ctx, int slot, const w80211_batch_in* in, w80211_batch_out* out)
{
    if (!ctx) return W80211_E_ARG;
    if (slot < 0 || slot >= ctx->nslots) return fail(ctx, W80211_E_ARG, "bad slot %d", slot);
    Slot& s = ctx->slots[slot];
    progress(ctx, -1);
    int r = enqueue_input_header(ctx, s, in, out, false);
    if (r < 0) {
        if (s.stream) cudaStreamSynchronize(s.stream);
        return r;
    }
    s.busy = true;
    s.phase = 1;
    s.cont_err = 0;
    ctx->pending.push_back(slot);
    return W80211_OK;
}

int w80211_llr_export(w80211_ctx* ctx, int slot, uint64_t* handle)
{
    if (!ctx || !handle) return W80211_E_ARG;
    *handle = 0;
    if (slot < 0 || slot >= ctx->nslots) return fail(ctx, W80211_E_ARG, "bad slot %d", slot);
    Slot& s = ctx->slots[slot];
    if (s.busy || !s.did_demod) return fail(ctx, W80211_E_STATE, "llr_export: slot %d holds no collected demod batch", slot);
    if (!s.llr.p || s.n_llr == 0) return fail(ctx, W80211_E_STATE, "llr_export: the batch in slot %d produced no LLRs", slot);
    std::lock_guard<std::mutex> lk(g_lease_mu);
    /* the newest W80211_LEASES_KEPT always stay; older ones stay too while the context's leases are few and small (a consumer
     * may lag by whatever the edge between the two blocks buffers) and retire oldest first beyond that (a decode batch still
     * reading one keeps its memory alive until it collects) */
    for (;;) {
        size_t held = s.llr.cap;
        for (uint64_t id : ctx->leases) {
            auto it = g_leases.find(id);
            if (it != g_leases.end()) held += it->second.buf.cap;
        }
        const size_t n = ctx->leases.size() + 1;
        if (n <= (size_t)W80211_LEASES_KEPT || (n <= 64 && held <= ctx->lease_bytes)) break;
        const uint64_t old = ctx->leases.front();
        ctx->leases.pop_front();
        auto it = g_leases.find(old);
        if (it == g_leases.end()) continue;
        it->second.retired = true;
        if (it->second.readers <= 0) lease_free_locked(old);
    }
    const uint64_t id = g_lease_next++;
    Lease& L = g_leases[id];
    L.device = ctx->device;
    L.buf = s.llr;
    L.nfloats = s.n_llr;
    L.owner = ctx;
    s.llr = DevBuf();
    s.n_llr = 0;
    ctx->leases.push_back(id);
    *handle = id;
    return W80211_OK;
}

int w80211_llr_query(uint64_t handle, int32_t* device, uint64_t* nfloats)
{
    std::lock_guard<std::mutex> lk(g_lease_mu);
    auto it = g_leases.find(handle);
    if (it == g_leases.end() || it->second.retired) return W80211_E_STATE;
    if (device) *device = it->second.device;
    if (nfloats) *nfloats = it->second.nfloats;
    return W80211_OK;
}

int w80211_lease_consumers(int delta)
{
    std::lock_guard<std::mutex> lk(g_lease_mu);
    g_lease_consumers = std::max(0, g_lease_consumers + delta);
    return g_lease_consumers;
}

int w80211_presiso_batch(w80211_ctx* ctx, const float* sig, uint64_t nsamples, int32_t mem, float* preac, float* preconj)
{
    NvtxRange nvtx_("w80211 presiso_batch");
    if (!ctx) return W80211_E_ARG;
    if (nsamples > 0 && (!sig || !preac || !preconj)) return fail(ctx, W80211_E_ARG, "bad presiso_batch arguments");
    if (nsamples == 0) return W80211_OK;
    CK(cudaSetDevice(ctx->device));
    SigStage& g = ctx->sig;
    if (!g.stream) {
        CK(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
        for (auto& e : g.ev) CK(cudaEventCreate(&e));
    }
    cudaStream_t st = g.stream;
    if (mem == W80211_MEM_DEVICE) {
        w_launch_presiso(st, (const float2*)sig, nsamples, preac, (float2*)preconj);
        CK(cudaGetLastError());
    } else {
        RS(dev_reserve(ctx, g.det_sig, (size_t)nsamples * 8 + 64));
        RS(dev_reserve(ctx, g.det_ac, (size_t)nsamples * 4 + 64));
        RS(dev_reserve(ctx, g.det_conj, (size_t)nsamples * 8 + 64));
        CK(cudaMemcpyAsync(g.det_sig.p, sig, (size_t)nsamples * 8, cudaMemcpyHostToDevice, st));
        w_launch_presiso(st, (const float2*)g.det_sig.p, nsamples, (float*)g.det_ac.p, (float2*)g.det_conj.p);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(preac, g.det_ac.p, (size_t)nsamples * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(preconj, g.det_conj.p, (size_t)nsamples * 8, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
    ctx->stats.kernel_launches += 1;
    return W80211_OK;
}

int w80211_detect_batch(w80211_ctx* ctx, const w80211_detect_in* in, w80211_detect_out* out)
{
    NvtxRange nvtx_("w80211 detect_batch");
    if (!ctx) return W80211_E_ARG;
    if (!in || !out || in->nsamples >= 0xffffffffull || (in->nsamples > 0 && (!in->preac || !in->preconj || !in->sig)) ||
        out->sync_capacity < 0 || (out->sync_capacity > 0 && !out->syncs))
        return fail(ctx, W80211_E_ARG, "bad detect_batch arguments");
    w80211_detect_state st0;
    memset(&st0, 0, sizeof(st0));
    if (in->state) st0 = *in->state;
    if (st0.n_plateau < 0 || st0.count_down < 0 || st0.count_down > 80 || st0.sync_skip < 0)
        return fail(ctx, W80211_E_ARG, "detect_batch: impossible carried state");
    const long long n = (long long)in->nsamples;
    static const bool trace = getenv("W80211_TRACE") != nullptr;
    auto tnow = [] { return std::chrono::steady_clock::now(); };
    auto tms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::milli>(b - a).count();
    };
    const auto tt0 = tnow();
    out->nsync = out->n_trigger = 0;
    out->next = in->nsamples;
    out->state = st0;
    if (n == 0) return W80211_OK;
    CK(cudaSetDevice(ctx->device));
    SigStage& g = ctx->sig;
    if (!g.stream) {
        CK(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
        for (auto& e : g.ev) CK(cudaEventCreate(&e));
    }
    cudaStream_t st = g.stream;
    const float* d_ac = in->preac;
    const float2* d_cj = (const float2*)in->preconj;
    const float2* d_sg = (const float2*)in->sig;
    if (in->mem != W80211_MEM_DEVICE) {
        RS(dev_reserve(ctx, g.det_ac, (size_t)n * 4 + 64));
        RS(dev_reserve(ctx, g.det_conj, (size_t)n * 8 + 64));
        RS(dev_reserve(ctx, g.det_sig, (size_t)n * 8 + 64));
        CK(cudaMemcpyAsync(g.det_ac.p, in->preac, (size_t)n * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(g.det_conj.p, in->preconj, (size_t)n * 8, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(g.det_sig.p, in->sig, (size_t)n * 8, cudaMemcpyHostToDevice, st));
        d_ac = (const float*)g.det_ac.p;
        d_cj = (const float2*)g.det_conj.p;
        d_sg = (const float2*)g.det_sig.p;
    }
    /* ---- runs of preac > 0.3: both edge lists in ascending order, counts and lists in one round trip.  The lists are sized
     * from what the previous call needed; a call that finds more edges than that grows them and runs the three small kernels
     * once more. */
    const int nblk = w_edges_blocks((unsigned long long)n);
    RS(dev_reserve(ctx, g.det_cnt, 16));
    RS(dev_reserve(ctx, g.det_tmp, (size_t)nblk * 2 * sizeof(uint2) + 16));
    unsigned int cap = std::max(g.edge_cap, 4096u), ns = 0, ne = 0;
    unsigned int *hs = nullptr, *he = nullptr;
    for (int attempt = 0;; attempt++) {
        RS(dev_reserve(ctx, g.det_starts, (size_t)cap * 4 + 16));
        RS(dev_reserve(ctx, g.det_ends, (size_t)cap * 4 + 16));
        RS(pin_reserve(ctx, g.h_edges, (size_t)cap * 8 + 16));
        CK(cudaMemsetAsync(g.det_cnt.p, 0, 16, st));
        w_launch_run_edges(st, d_ac, (unsigned long long)n, st0.n_plateau > 0, cap, (unsigned int*)g.det_cnt.p, (unsigned int*)g.det_starts.p,
                           (unsigned int*)g.det_ends.p, g.det_tmp.p);
        CK(cudaGetLastError());
        unsigned int* hc = (unsigned int*)g.h_edges.p;
        hs = hc + 4;
        he = hs + cap;
        CK(cudaMemcpyAsync(hc, g.det_cnt.p, 16, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(hs, g.det_starts.p, (size_t)cap * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(he, g.det_ends.p, (size_t)cap * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        ns = hc[0];
        ne = hc[1];
        if (std::max(ns, ne) <= cap) break;
        if (attempt) return fail(ctx, W80211_E_STATE, "detect_batch: edge count changed between two passes");
        cap = std::max(ns, ne) + 16;
    }
    g.edge_cap = std::max(ns, ne) + std::max(ns, ne) / 2 + 16;
    const auto tt1 = tnow();
    struct Run {
        long long s, e; /* [s, e): s < 0 when the run began before this buffer, e = n when it is still open */
        long long i0;   /* sample at which this run started a countdown, -1: none */
        bool fpe0;      /* carried d_fPlateauEnd */
    };
    std::vector<Run> runs;
    runs.reserve((size_t)ns + 1);
    const bool cont = st0.n_plateau > 0;
    if (cont) runs.push_back(Run{ -(long long)st0.n_plateau, n, -1, st0.f_plateau_end != 0 });
    for (unsigned int k = 0; k < ns; k++) runs.push_back(Run{ (long long)hs[k], n, -1, false });
    if (ne > runs.size() || ne + 1 < runs.size()) return fail(ctx, W80211_E_STATE, "detect_batch: run edges do not pair up");
    for (unsigned int k = 0; k < ne; k++) runs[k].e = (long long)he[k];
    /* ---- the trigger block's plateau / countdown logic over the runs (trigger_impl.cc:76-108) */
    std::vector<long long> trig; /* samples that carry the 0x01 flag, ascending; may lie beyond this buffer */
    long long busy_until = -1;   /* d_fPlateau is set up to and including this sample */
    if (st0.count_down > 0) {
        busy_until = (long long)st0.count_down - 1;
        trig.push_back(busy_until);
    }
    for (auto& r : runs) {
        if (r.fpe0) continue; /* one countdown per run */
        const long long i0 = std::max(std::max(r.s + 20, busy_until + 1), 0ll);
        if (i0 < r.e) {
            r.i0 = i0;
            busy_until = i0 + 79;
            trig.push_back(busy_until);
        }
    }
    /* ---- the sync block's IDLE / SYNC alternation (sync_impl.cc:73-134) */
    struct Seen {
        long long t, pos;
    };
    std::vector<Seen> seen;
    long long pos = st0.sync_skip, next = n;
    for (long long t : trig) {
        if (t >= n) break;
        if (t < pos) continue; /* inside the 111 samples the previous SYNC pass consumed */
        if (t + 240 > n) {
            if (!in->final) next = t;
            break;
        }
        seen.push_back(Seen{ t, pos });
        pos = t + 111;
    }
    if ((long long)seen.size() > (long long)out->sync_capacity)
        return fail(ctx, W80211_E_CAPACITY, "detect_batch: sync_capacity %d < %zu triggers", out->sync_capacity, seen.size());
    /* ---- per-trigger work on the GPU, plus one more record for the state at `next` */
    const int nev = (int)seen.size() + 1;
    RS(pin_reserve(ctx, g.h_ev, (size_t)nev * sizeof(w_det_event)));
    RS(pin_reserve(ctx, g.h_dres, (size_t)nev * sizeof(w_det_result)));
    RS(dev_reserve(ctx, g.det_ev, (size_t)nev * sizeof(w_det_event)));
    RS(dev_reserve(ctx, g.det_res, (size_t)nev * sizeof(w_det_result)));
    w_det_event* hev = (w_det_event*)g.h_ev.p;
    size_t ri = 0; /* last run that starts before the event's sample */
    auto last_run_before = [&](long long t) -> const Run* {
        while (ri + 1 < runs.size() && runs[ri + 1].s < t) ri++;
        return (!runs.empty() && runs[ri].s < t) ? &runs[ri] : nullptr;
    };
    for (int k = 0; k < nev; k++) {
        const bool real = k + 1 < nev;
        const long long t = real ? seen[(size_t)k].t : next;
        const Run* r = last_run_before(t);
        hev[k].rs = r ? r->s : 0;
        hev[k].re = r ? std::min(r->e, t) : 0;
        hev[k].pos = real ? seen[(size_t)k].pos : pos;
        hev[k].t = t;
        hev[k].do_sync = real ? 1 : 0;
        hev[k].pad_ = 0;
    }
    const auto tt2 = tnow();
    CK(cudaMemcpyAsync(g.det_ev.p, hev, (size_t)nev * sizeof(w_det_event), cudaMemcpyHostToDevice, st));
    w_launch_sync_events(st, nev, (const w_det_event*)g.det_ev.p, d_ac, d_cj, d_sg, st0.conj_ac, make_float2(st0.conj_re, st0.conj_im),
                         (w_det_result*)g.det_res.p);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(g.h_dres.p, g.det_res.p, (size_t)nev * sizeof(w_det_result), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    const auto tt3 = tnow();
    if (trace)
        fprintf(stderr, "detect_batch: edges %.2f ms (%u starts), host walk %.2f ms, events %.2f ms (%d)\n", tms(tt0, tt1), ns, tms(tt1, tt2),
                tms(tt2, tt3), nev);
    const w_det_result* hr = (const w_det_result*)g.h_dres.p;
    int nsync = 0;
    for (size_t k = 0; k < seen.size(); k++) {
        if (!hr[k].valid) continue;
        w80211_sync& sy = out->syncs[nsync++];
        sy.pos = (uint64_t)(seen[k].t + hr[k].mindex);
        sy.rad = hr[k].rad;
        sy.snr = hr[k].snr;
        sy.rssi = hr[k].rssi;
        sy.flags = W80211_SYNC_TAGGED;
    }
    out->nsync = nsync;
    out->n_trigger = (int32_t)seen.size();
    out->next = (uint64_t)next;
    /* ---- both blocks' state at `next` */
    w80211_detect_state s1;
    memset(&s1, 0, sizeof(s1));
    const w_det_result& tail = hr[nev - 1];
    s1.conj_re = tail.conj.x;
    s1.conj_im = tail.conj.y;
    for (const auto& r : runs) {
        if (r.s < next && r.e >= next && next > 0) { /* sample next-1 is inside this run */
            s1.n_plateau = (int32_t)std::min<long long>(next - r.s, 0x7fffffff);
            s1.f_plateau_end = (r.fpe0 || (r.i0 >= 0 && r.i0 < next)) ? 1 : 0;
            s1.conj_ac = tail.run_max;
            break;
        }
    }
    for (long long t : trig) {
        if (t < next) continue;
        if (t - 79 < next) s1.count_down = (int32_t)(t - next + 1);
        break;
    }
    s1.sync_skip = (int32_t)std::max(0ll, pos - next);
    out->state = next > 0 ? s1 : st0; /* next = 0: nothing was consumed */
    ctx->stats.kernel_launches += 4;
    return W80211_OK;
}

int w80211_detect_stream(w80211_ctx* ctx, const float* sig, uint64_t nsamples, uint64_t lead, int32_t mem, int32_t final,
                         const w80211_detect_state* state, w80211_detect_out* out)
{
    NvtxRange nvtx_("w80211 detect_stream");
    if (!ctx) return W80211_E_ARG;
    if (!out || lead > nsamples || nsamples >= 0xffffffffull || (nsamples > 0 && !sig)) return fail(ctx, W80211_E_ARG, "bad detect_stream arguments");
    CK(cudaSetDevice(ctx->device));
    SigStage& g = ctx->sig;
    if (!g.stream) {
        CK(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
        for (auto& e : g.ev) CK(cudaEventCreate(&e));
    }
    cudaStream_t st = g.stream;
    const float2* d_sg = (const float2*)sig;
    if (nsamples > 0) {
        if (mem != W80211_MEM_DEVICE) {
            RS(dev_reserve(ctx, g.str_sig, (size_t)nsamples * 8 + 64));
            CK(cudaMemcpyAsync(g.str_sig.p, sig, (size_t)nsamples * 8, cudaMemcpyHostToDevice, st));
            d_sg = (const float2*)g.str_sig.p;
        }
        RS(dev_reserve(ctx, g.str_ac, (size_t)nsamples * 4 + 64));
        RS(dev_reserve(ctx, g.str_conj, (size_t)nsamples * 8 + 64));
        w_launch_presiso(st, d_sg, nsamples, (float*)g.str_ac.p, (float2*)g.str_conj.p);
        CK(cudaGetLastError());
        ctx->stats.kernel_launches += 1;
    }
    /* the detection proper, on the same stream, from the device-resident streams past the history */
    w80211_detect_in din;
    memset(&din, 0, sizeof(din));
    din.mem = W80211_MEM_DEVICE;
    din.final = final;
    din.nsamples = nsamples - lead;
    din.preac = (const float*)g.str_ac.p + lead;
    din.preconj = (const float*)((const float2*)g.str_conj.p + lead);
    din.sig = (const float*)(d_sg + lead);
    din.state = state;
    return w80211_detect_batch(ctx, &din, out);
}

int w80211_signal_batch(w80211_ctx* ctx, const w80211_signal_in* in, w80211_signal_out* out)
{
    NvtxRange nvtx_("w80211 signal_batch");
    if (!ctx) return W80211_E_ARG;
    if (!in || !out || in->nsync < 0 || (in->nant != 1 && in->nant != 2) || (in->nsync > 0 && !in->syncs) ||
        (in->nsamples > 0 && (!in->iq[0] || (in->nant == 2 && !in->iq[1]))) || in->start > in->nsamples)
        return fail(ctx, W80211_E_ARG, "bad signal_batch arguments");
    if (in->nsync > 0 && (!out->pkts || !out->chan)) return fail(ctx, W80211_E_ARG, "signal_batch: pkts / chan missing");
    CK(cudaSetDevice(ctx->device));
    SigStage& g = ctx->sig;
    if (!g.stream) {
        CK(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
        for (auto& e : g.ev) CK(cudaEventCreate(&e));
    }
    cudaStream_t st = g.stream;
    const int ns = in->nsync, nant = in->nant;
    out->lsig_ms = out->copy_ms = 0.f;
    out->npkt = out->nsync_used = out->n_rejected = 0;
    out->iq_used = 0;
    out->next = in->start;
    out->need = 0;
    for (int i = 1; i < ns; i++)
        if (in->syncs[i].pos < in->syncs[i - 1].pos) return fail(ctx, W80211_E_ARG, "signal_batch: syncs not in ascending order");
    if (ns == 0) {
        out->next = in->nsamples;
        return W80211_OK;
    }
    /* raw samples */
    const float2* raw[2] = {nullptr, nullptr};
    for (int a = 0; a < nant; a++) {
        if (in->mem == W80211_MEM_DEVICE) {
            raw[a] = (const float2*)in->iq[a];
        } else {
            RS(dev_reserve(ctx, g.iq[a], (size_t)in->nsamples * 8 + 64));
            CK(cudaMemcpyAsync(g.iq[a].p, in->iq[a], (size_t)in->nsamples * 8, cudaMemcpyHostToDevice, st));
            raw[a] = (const float2*)g.iq[a].p;
        }
    }
    /* every flag's L-SIG at once */
    RS(dev_reserve(ctx, g.syncs, (size_t)ns * sizeof(w80211_sync)));
    RS(dev_reserve(ctx, g.res, (size_t)ns * sizeof(int4)));
    RS(dev_reserve(ctx, g.chan_all, (size_t)ns * 64 * 8));
    RS(pin_reserve(ctx, g.h_res, (size_t)ns * sizeof(int4)));
    CK(cudaMemcpyAsync(g.syncs.p, in->syncs, (size_t)ns * sizeof(w80211_sync), cudaMemcpyHostToDevice, st));
    CK(cudaEventRecord(g.ev[0], st));
    w_launch_lsig(st, ctx->d_tab, ns, (const w80211_sync*)g.syncs.p, raw[0], in->nsamples, (int4*)g.res.p, (float2*)g.chan_all.p);
    CK(cudaGetLastError());
    CK(cudaEventRecord(g.ev[1], st));
    CK(cudaMemcpyAsync(g.h_res.p, g.res.p, (size_t)ns * sizeof(int4), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    /* the block's state machine over the flags, in stream order (signal_impl.cc:76-160,194-201) */
    const int4* hr = (const int4*)g.h_res.p;
    RS(pin_reserve(ctx, g.h_acc, (size_t)ns * sizeof(w_sig_acc)));
    w_sig_acc* acc = (w_sig_acc*)g.h_acc.p;
    uint64_t resume = in->start, used = 0, need = 0;
    int np = 0, max_nsamp = 0, i = 0;
    bool stopped = false;
    for (; i < ns; i++) {
        const w80211_sync& sy = in->syncs[i];
        if (sy.pos < resume) continue; /* consumed inside a packet / the 80 samples after a failed check */
        if (!(sy.flags & W80211_SYNC_TAGGED)) {
            resume = sy.pos + 1;
            continue;
        }
        if (sy.pos + 224 > in->nsamples) {
            stopped = true;
            need = sy.pos + 224;
            break;
        }
        if (hr[i].x != 1) {
            out->n_rejected++;
            resume = sy.pos + 80;
            continue;
        }
        const int nsamp = hr[i].w;
        if (sy.pos + 224 + (uint64_t)nsamp > in->nsamples) {
            stopped = true;
            need = sy.pos + 224 + (uint64_t)nsamp;
            break;
        }
        if (used + (uint64_t)nsamp > out->iq_capacity)
            return fail(ctx, W80211_E_ARG, "signal_batch: iq_capacity %llu too small", (unsigned long long)out->iq_capacity);
        w80211_pkt_in& pk = out->pkts[np];
        pk.iq_offset = used;
        pk.nsamp = nsamp;
        pk.l_mcs = hr[i].y;
        pk.l_len = hr[i].z;
        pk.cfo = sy.rad * 3183098.8618379068f; /* rad * 20e6 / 2pi, binary32 like signal_impl.cc:136 */
        pk.snr = sy.snr;
        pk.rssi = sy.rssi;
        pk.iq_scale = 1.0f;
        if (out->sync_of_pkt) out->sync_of_pkt[np] = i;
        acc[np].src = sy.pos;
        acc[np].dst = used;
        acc[np].nsamp = nsamp;
        acc[np].rad = sy.rad;
        acc[np].sync = i;
        acc[np].pad_ = 0;
        max_nsamp = std::max(max_nsamp, nsamp);
        used += (uint64_t)nsamp;
        np++;
        resume = sy.pos + 224 + (uint64_t)nsamp;
    }
    out->npkt = np;
    out->iq_used = used;
    if (stopped && !in->final) {
        out->nsync_used = i;
        out->next = in->syncs[i].pos;
        out->need = need;
    } else {
        out->nsync_used = ns;
        out->next = in->nsamples;
    }
    CK(cudaEventElapsedTime(&out->lsig_ms, g.ev[0], g.ev[1]));
    if (np == 0) return W80211_OK;
    /* copy the accepted packets with their rotation */
    for (int a = 0; a < nant; a++)
        if (!out->iq[a]) return fail(ctx, W80211_E_ARG, "signal_batch: output iq[%d] missing", a);
    float2* dst[2] = {nullptr, nullptr};
    float2* dchan = nullptr;
    if (out->mem == W80211_MEM_DEVICE) {
        dst[0] = (float2*)out->iq[0];
        dst[1] = (float2*)out->iq[1];
        dchan = (float2*)out->chan;
    } else {
        for (int a = 0; a < nant; a++) {
            RS(dev_reserve(ctx, g.out_iq[a], (size_t)used * 8 + 64));
            dst[a] = (float2*)g.out_iq[a].p;
        }
        RS(dev_reserve(ctx, g.out_chan, (size_t)np * 64 * 8));
        dchan = (float2*)g.out_chan.p;
    }
    RS(dev_reserve(ctx, g.acc, (size_t)np * sizeof(w_sig_acc)));
    CK(cudaMemcpyAsync(g.acc.p, acc, (size_t)np * sizeof(w_sig_acc), cudaMemcpyHostToDevice, st));
    CK(cudaEventRecord(g.ev[2], st));
    w_launch_cfo_copy(st, np, nant, max_nsamp, (const w_sig_acc*)g.acc.p, raw[0], raw[1], dst[0], dst[1],
                      (const float2*)g.chan_all.p, dchan);
    CK(cudaGetLastError());
    CK(cudaEventRecord(g.ev[3], st));
    if (out->mem != W80211_MEM_DEVICE) {
        for (int a = 0; a < nant; a++) CK(cudaMemcpyAsync(out->iq[a], dst[a], (size_t)used * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(out->chan, dchan, (size_t)np * 64 * 8, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
    CK(cudaEventElapsedTime(&out->copy_ms, g.ev[2], g.ev[3]));
    ctx->stats.kernel_launches += 2;
    return W80211_OK;
}

int w80211_demod_batch(w80211_ctx* ctx, const w80211_batch_in* in, w80211_batch_out* out)
{
    if (!ctx) return W80211_E_ARG;
    Slot& s = ctx->slots[0];
    progress(ctx, -2);
    RS(enqueue_input_header(ctx, s, in, out, false));
    s.busy = true;
    int r = cudaEventSynchronize(s.ev[EV_HDRD]) == cudaSuccess ? 0 : fail(ctx, W80211_E_CUDA, "header stage failed");
    if (r == 0) r = continue_front(ctx, s);
    if (r == 0) r = finish_outputs(ctx, s);
    if (r < 0) {
        cudaStreamSynchronize(s.stream);
        s.busy = false;
        return r;
    }
    return collect_slot(ctx, s);
}

/* LLRs + descriptors -> everything of the decode half enqueued on the slot's stream; does not wait */
static int decode_submit_impl(w80211_ctx* ctx, int slot, const w80211_decode_in* in, w80211_batch_out* out)
{
    NvtxRange nvtx_("w80211 decode_submit");
    if (slot < 0 || slot >= ctx->nslots) return fail(ctx, W80211_E_ARG, "bad slot %d", slot);
    if (!in || !out || in->npkt < 0 || (in->npkt > 0 && !in->desc) || in->llr_mem < W80211_MEM_HOST || in->llr_mem > W80211_MEM_LEASE ||
        (in->npkt > 0 && in->llr_mem != W80211_MEM_LEASE && !in->llr))
        return fail(ctx, W80211_E_ARG, "bad decode arguments");
    Slot& s = ctx->slots[slot];
    if (s.busy) return fail(ctx, W80211_E_STATE, "slot busy: collect it first");
    CK(cudaSetDevice(ctx->device));
    progress(ctx, -1);
    cudaStream_t st = s.stream;
    const int npkt = in->npkt;
    const int llr_mem = in->llr_mem;
    const float* llr = in->llr;
    uint64_t nllr = in->nllr;
    s.npkt = npkt;
    s.out = out;
    s.did_demod = false;
    s.did_decode = true;
    s.in_host = llr_mem == W80211_MEM_HOST;
    s.launches = 0;
    s.llr_src = nullptr;
    s.lease = 0;
    out->n_mpdu = out->n_crc_ok = 0;
    out->n_bytes = out->n_llr = out->psdu_bits_ok = 0;
    if (llr_mem == W80211_MEM_LEASE) {
        std::lock_guard<std::mutex> lk(g_lease_mu);
        auto it = g_leases.find(in->lease);
        if (it == g_leases.end() || it->second.retired) return fail(ctx, W80211_E_STATE, "decode: LLR lease %llu has been retired", (unsigned long long)in->lease);
        if (it->second.device != ctx->device)
            return fail(ctx, W80211_E_ARG, "decode: LLR lease lives on device %d, this context on %d", it->second.device, ctx->device);
        it->second.readers++;
        s.lease = in->lease;
        llr = (const float*)it->second.buf.p;
        nllr = it->second.nfloats;
    }
    auto bail = [&](int r) {
        lease_release(s.lease);
        s.lease = 0;
        return r;
    };
#define CKB(call)                                                                                                       \
    do {                                                                                                                \
        cudaError_t e_ = (call);                                                                                        \
        if (e_ != cudaSuccess) return bail(fail(ctx, W80211_E_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_))); \
    } while (0)
#define RSB(call)                  \
    do {                           \
        int r_ = (call);           \
        if (r_ < 0) return bail(r_); \
    } while (0)
    CKB(cudaEventRecord(s.ev[EV_START], st));
    RSB(pin_reserve(ctx, s.h_desc, (size_t)npkt * sizeof(w80211_desc) + 16));
    RSB(dev_reserve(ctx, s.desc, (size_t)npkt * sizeof(w80211_desc) + 16));
    w80211_desc* hd = (w80211_desc*)s.h_desc.p;
    if (llr_mem == W80211_MEM_DEVICE) {
        if (npkt) CKB(cudaMemcpy(hd, in->desc, (size_t)npkt * sizeof(w80211_desc), cudaMemcpyDeviceToHost));
    } else if (npkt) {
        memcpy(hd, in->desc, (size_t)npkt * sizeof(w80211_desc));
    }
    std::vector<uint64_t>& src_off = s.v_src_off;
    src_off.resize((size_t)npkt);
    /* The kernels need every packet's LLRs at a 16-byte aligned offset.  A caller that already laid them out that way
     * (gr/lib/decode_impl.cc, a lease, the LLR stream demod_batch returned) is decoded from where they are -- device
     * memory and leases without any copy, host memory with one copy of the whole range; anything else is repacked packet
     * by packet into the slot's buffer. */
    bool in_place = true;
    uint64_t lo = ~0ull, hi = 0;
    for (int p = 0; p < npkt; p++) {
        w80211_desc& d = hd[p];
        src_off[(size_t)p] = (uint64_t)d.llr_off;
        const bool ok = d.status == W80211_ST_OK && d.total > 0 && d.llr_off >= 0 && (uint64_t)d.llr_off + (uint64_t)d.total <= nllr;
        if (!ok) {
            if (d.status == W80211_ST_OK) d.status = W80211_ST_BAD_INPUT;
            d.llr_off = -1;
            continue;
        }
        if (d.llr_off & 3) in_place = false;
        lo = std::min(lo, (uint64_t)d.llr_off);
        hi = std::max(hi, (uint64_t)d.llr_off + (uint64_t)d.total);
    }
    if (((uintptr_t)llr & 15u) != 0) in_place = false;
    if (llr_mem == W80211_MEM_LEASE && !in_place) return bail(fail(ctx, W80211_E_ARG, "decode: lease offsets must be multiples of 4"));
    if (lo > hi) lo = hi = 0;
    if (in_place && llr_mem != W80211_MEM_HOST) {
        s.llr_src = llr;
        s.n_llr = 0;
    } else if (in_place) {
        lo &= ~(uint64_t)3;
        RSB(llr_reserve(ctx, s, (size_t)(hi - lo + 64) * 4));
        if (hi > lo) CKB(cudaMemcpyAsync(s.llr.p, llr + lo, (size_t)(hi - lo) * 4, cudaMemcpyHostToDevice, st));
        for (int p = 0; p < npkt; p++)
            if (hd[p].llr_off >= 0) hd[p].llr_off -= (int64_t)lo;
        s.n_llr = hi - lo;
    } else {
        uint64_t off = 0;
        for (int p = 0; p < npkt; p++) {
            w80211_desc& d = hd[p];
            if (d.llr_off < 0) continue;
            d.llr_off = (int64_t)off;
            off += align_up((size_t)d.total, 4) + 16;
        }
        s.n_llr = off;
        RSB(llr_reserve(ctx, s, (size_t)(off + 64) * 4));
        for (int p = 0; p < npkt; p++) {
            const w80211_desc& d = hd[p];
            if (d.llr_off < 0) continue;
            CKB(cudaMemcpyAsync((float*)s.llr.p + d.llr_off, llr + src_off[(size_t)p], (size_t)d.total * 4,
                                llr_mem == W80211_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
        }
    }
    CKB(cudaEventRecord(s.ev[EV_H2D], st));
    CKB(cudaEventRecord(s.ev[EV_HDR], st));
    const auto tp0 = std::chrono::steady_clock::now();
    RSB(plan_and_upload(ctx, s, hd, false, true));
    s.host_plan_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - tp0).count();
    CKB(cudaEventRecord(s.ev[EV_PLAN], st));
    CKB(cudaEventRecord(s.ev[EV_DEMOD], st));
    int r = enqueue_decode_tail(ctx, s);
    if (r == 0) {
        /* the descriptors go back with the caller's own LLR offsets (decode_batch, after collect) */
        const bool want = out->want_llr;
        w80211_desc* user_desc = out->desc;
        out->want_llr = 0;
        out->desc = nullptr;
        r = finish_outputs(ctx, s);
        out->want_llr = want;
        out->desc = user_desc;
    }
    if (r < 0) {
        cudaStreamSynchronize(s.stream);
        return bail(r);
    }
#undef CKB
#undef RSB
    s.busy = true;
    s.phase = 2;
    s.cont_err = 0;
    return W80211_OK;
}

int w80211_decode_submit(w80211_ctx* ctx, int slot, const w80211_decode_in* in, w80211_batch_out* out)
{
    if (!ctx) return W80211_E_ARG;
    return decode_submit_impl(ctx, slot, in, out);
}

int w80211_decode_batch(w80211_ctx* ctx, int32_t npkt, const w80211_desc* desc, const float* llr, uint64_t nllr,
                        int32_t llr_mem, w80211_batch_out* out)
{
    if (!ctx) return W80211_E_ARG;
    if (llr_mem != W80211_MEM_HOST && llr_mem != W80211_MEM_DEVICE) return fail(ctx, W80211_E_ARG, "bad decode_batch arguments");
    progress(ctx, -2);
    w80211_decode_in in;
    memset(&in, 0, sizeof(in));
    in.npkt = npkt;
    in.llr_mem = llr_mem;
    in.desc = desc;
    in.llr = llr;
    in.nllr = nllr;
    int r = decode_submit_impl(ctx, 0, &in, out);
    if (r < 0) return r;
    r = w80211_rx_collect(ctx, 0);
    if (r == 0 && out->desc && out->mem == W80211_MEM_HOST) {
        const Slot& s = ctx->slots[0];
        const w80211_desc* hd = (const w80211_desc*)s.h_desc.p;
        for (int p = 0; p < npkt; p++) {
            out->desc[p] = hd[p];
            out->desc[p].llr_off = (int64_t)s.v_src_off[(size_t)p];
        }
    }
    return r;
}

} /* extern "C" */
