/* demod_core.cc -- packet gathering, non-blocking GPU batches and LLR emission shared by demod and demod2.
 * See demod_core.h for the structure; reference behaviour is cited where it is mirrored. */
#include "demod_core.h"

#include <algorithm>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#if defined(__x86_64__) && defined(__GNUC__)
#include <immintrin.h>
#endif

namespace gr {
namespace ieee80211 {

/* ------------------------------------------------------------------ staged_array */
template <class T>
bool staged_array<T>::reserve(size_t n, size_t keep)
{
    if (n <= d_cap) return true;
    size_t want = std::max(n + n / 2, (size_t)4096);
    bool pinned = true;
    T* np = static_cast<T*>(w80211_host_alloc(want * sizeof(T)));
    if (!np) {
        pinned = false;
        np = static_cast<T*>(std::malloc(want * sizeof(T)));
        if (!np) return false;
    }
    if (d_p && keep) std::memcpy(np, d_p, std::min(keep, d_cap) * sizeof(T));
    release();
    d_p = np;
    d_cap = want;
    d_pinned = pinned;
    return true;
}

template <class T>
void staged_array<T>::release()
{
    if (d_p) {
        if (d_pinned)
            w80211_host_free(d_p);
        else
            std::free(d_p);
    }
    d_p = nullptr;
    d_cap = 0;
}

template class staged_array<float>;
template class staged_array<unsigned char>;

/* ------------------------------------------------------------------ stream_copy */
/* Samples go from the scheduler's buffer into page-locked staging exactly once and are next read by the GPU's copy engine, not
 * by this core: written with non-temporal stores they neither cost a read-for-ownership of the destination line nor push the
 * source out of the cache.  AVX2 when the CPU has it (checked at run time), memcpy otherwise. */
#if defined(__x86_64__) && defined(__GNUC__)
__attribute__((target("avx2"))) static void stream_copy_avx2(unsigned char* d, const unsigned char* s, size_t n)
{
    const size_t head = (32 - (reinterpret_cast<uintptr_t>(d) & 31)) & 31;
    if (head >= n) {
        std::memcpy(d, s, n);
        return;
    }
    std::memcpy(d, s, head);
    d += head, s += head, n -= head;
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        const __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(s + i));
        const __m256i b = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(s + i + 32));
        const __m256i c = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(s + i + 64));
        const __m256i e = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(s + i + 96));
        _mm256_stream_si256(reinterpret_cast<__m256i*>(d + i), a);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(d + i + 32), b);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(d + i + 64), c);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(d + i + 96), e);
    }
    _mm_sfence();
    if (i < n) std::memcpy(d + i, s + i, n - i);
}
void stream_copy(void* dst, const void* src, size_t bytes)
{
    static const bool avx2 = __builtin_cpu_supports("avx2") && std::getenv("W80211_NO_STREAM_COPY") == nullptr;
    if (avx2 && bytes >= 4096)
        stream_copy_avx2(static_cast<unsigned char*>(dst), static_cast<const unsigned char*>(src), bytes);
    else
        std::memcpy(dst, src, bytes);
}
#else
void stream_copy(void* dst, const void* src, size_t bytes) { std::memcpy(dst, src, bytes); }
#endif

/* ------------------------------------------------------------------ rx_env */
static long env_long(const char* name, long dflt)
{
    const char* v = std::getenv(name);
    if (!v || !*v) return dflt;
    char* end = nullptr;
    long r = std::strtol(v, &end, 10);
    return end && end != v ? r : dflt;
}

rx_env rx_env::read()
{
    rx_env e;
    e.device = (int)env_long("W80211_DEVICE", 0);
    e.max_batch = (int)std::max(1L, env_long("W80211_MAX_BATCH", 2048));
    e.min_batch = (int)std::max(1L, env_long("W80211_MIN_BATCH", 1));
    const char* w = std::getenv("W80211_MAX_WAIT_MS");
    e.max_wait_ms = w && *w ? std::atof(w) : 5.0;
    e.handoff = HANDOFF_AUTO;
    if (const char* h = std::getenv("W80211_LLR_HANDOFF")) {
        const std::string s(h);
        if (s == "host")
            e.handoff = HANDOFF_HOST;
        else if (s == "both")
            e.handoff = HANDOFF_BOTH;
        else if (s == "device")
            e.handoff = HANDOFF_DEVICE;
        else if (!s.empty() && s != "auto")
            std::fprintf(stderr, "ieee80211: W80211_LLR_HANDOFF=%s not understood (auto | host | both | device), using auto\n", h);
    }
    return e;
}

/* ------------------------------------------------------------------ demod_core */
demod_core::demod_core(gr::block* owner, int nant, const char* label, int mupos)
    : d_blk(owner),
      d_nant(nant),
      d_mupos(mupos ? 1 : 0),
      d_label(label),
      d_env(rx_env::read()),
      d_ctx(nullptr),
      d_failed(false),
      d_cur(-1),
      d_partial_in(-1),
      d_left(0),
      d_copy_left(0),
      d_trace(std::getenv("W80211_TRACE") != nullptr),
      d_t_ingest(0),
      d_t_submit(0),
      d_t_wait(0),
      d_t_emit(0),
      d_n_work(0),
      d_n_batches(0),
      d_n_pkts(0)
{
    for (bool& b : d_slot_busy) b = false;
    w80211_cfg cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.device = d_env.device;
    cfg.max_packets = d_env.max_batch;
    cfg.nslots = NSLOTS;
    int r = w80211_abi_version() == W80211_ABI_VERSION ? w80211_create(&cfg, &d_ctx) : W80211_E_ARG;
    if (w80211_abi_version() != W80211_ABI_VERSION)
        std::fprintf(stderr, "ieee80211 %s: libw80211rx.so has ABI version %d, this block was built for %d\n", d_label.c_str(),
                     w80211_abi_version(), W80211_ABI_VERSION);
    if (r != W80211_OK) {
        d_ctx = nullptr;
        d_failed = true;
        std::fprintf(stderr,
                     "ieee80211 %s: w80211_create failed (%d): no usable CUDA device %d; there is no CPU path, "
                     "packets will be dropped\n",
                     d_label.c_str(), r, d_env.device);
    }
    take_free_set();
}

demod_core::~demod_core()
{
    if (d_trace)
        std::fprintf(stderr, "ieee80211 %s trace: %llu general_work calls, %llu batches, %llu packets; ingest %.1f ms, submit %.1f ms, "
                             "waiting for the GPU %.1f ms, emit %.1f ms\n",
                     d_label.c_str(), (unsigned long long)d_n_work, (unsigned long long)d_n_batches, (unsigned long long)d_n_pkts,
                     1e3 * d_t_ingest, 1e3 * d_t_submit, 1e3 * d_t_wait, 1e3 * d_t_emit);
    if (d_ctx) {
        for (int idx : d_queue)
            if (d_sets[idx].state == SET_INFLIGHT) w80211_rx_collect(d_ctx, d_sets[idx].slot);
        w80211_destroy(d_ctx);
    }
}

void demod_core::report(const char* what)
{
    std::fprintf(stderr, "ieee80211 %s: %s: %s; batch dropped\n", d_label.c_str(), what, w80211_last_error(d_ctx));
}

bool demod_core::flush_pending() const { return !d_queue.empty() || (d_cur >= 0 && d_sets[d_cur].ncomplete > 0); }

void demod_core::forecast(int /*noutput_items*/, gr_vector_int& ninput_items_required)
{
    /* unfinished batches or whole staged packets can be turned into output without new input */
    const int need = flush_pending() ? 0 : 1;
    for (auto& n : ninput_items_required) n = need;
}

/* A free set becomes the one being filled; the packet still arriving when the previous one was submitted moves over. */
bool demod_core::take_free_set()
{
    if (d_cur >= 0) return true;
    int idx = -1;
    for (int i = 0; i < NSETS; i++)
        if (d_sets[i].state == SET_FREE) {
            idx = i;
            break;
        }
    if (idx < 0) return false;
    batch_set& b = d_sets[idx];
    b.state = SET_FILLING;
    b.pkts.clear();
    b.fill = 0;
    b.ncomplete = 0;
    if (d_partial_in >= 0) {
        batch_set& src = d_sets[d_partial_in];
        w80211_pkt_in last = src.pkts.back();
        const uint64_t have = src.fill - last.iq_offset;
        bool ok = b.chan.reserve(128, 0);
        for (int a = 0; a < d_nant; a++) ok = ok && b.iq[a].reserve((size_t)(have + d_copy_left + 16) * 2, 0);
        if (ok) {
            for (int a = 0; a < d_nant; a++)
                if (have) std::memcpy(b.iq[a].data(), src.iq[a].data() + (size_t)last.iq_offset * 2, (size_t)have * 8);
            std::memcpy(b.chan.data(), src.chan.data() + (size_t)src.npkt * 128, 128 * sizeof(float));
        } else {
            if (!d_failed) std::fprintf(stderr, "ieee80211 %s: out of host memory for staging, dropping packets\n", d_label.c_str());
            d_failed = true;
        }
        last.iq_offset = 0;
        b.pkts.push_back(last);
        b.fill = have;
        d_partial_in = -1;
    }
    d_cur = idx;
    return true;
}

/* Tags on the first item of a packet, keys and defaults as at demod_impl.cc:70-89. */
bool demod_core::start_packet(uint64_t abs_offset)
{
    std::vector<gr::tag_t> tags;
    d_blk->get_tags_in_range(tags, 0, abs_offset, abs_offset + 1);
    if (tags.empty()) return false;
    batch_set& b = d_sets[d_cur];
    w80211_pkt_in pk;
    std::memset(&pk, 0, sizeof(pk));
    long nsamp = -1;
    pk.l_mcs = -1;
    pk.l_len = -1;
    std::vector<std::complex<float>> chan;
    for (const auto& t : tags) {
        const std::string key = pmt::symbol_to_string(t.key);
        if (key == "cfo")
            pk.cfo = (float)pmt::to_float(t.value);
        else if (key == "snr")
            pk.snr = (float)pmt::to_float(t.value);
        else if (key == "rssi")
            pk.rssi = (float)pmt::to_float(t.value);
        else if (key == "mcs")
            pk.l_mcs = (int32_t)pmt::to_long(t.value);
        else if (key == "len")
            pk.l_len = (int32_t)pmt::to_long(t.value);
        else if (key == "nsamp")
            nsamp = pmt::to_long(t.value);
        else if (key == "chan")
            chan = pmt::c32vector_elements(t.value);
        /* "seq" is only printed by the reference */
    }
    if (nsamp + 320 <= 0 || nsamp > 0x3fffffff) {
        /* a tag no signal block writes: the tagged item is passed over (the reference would wait on it for ever) */
        d_left = 1;
        d_copy_left = 0;
        return true;
    }
    d_left = nsamp + 320; /* demod_impl.cc:92 */
    d_copy_left = std::max(nsamp, 0L);
    pk.nsamp = (int32_t)d_copy_left;
    pk.iq_offset = b.fill;
    const size_t np = b.pkts.size();
    bool ok = b.chan.reserve((np + 1) * 128, np * 128);
    for (int a = 0; a < d_nant; a++) ok = ok && b.iq[a].reserve((size_t)(b.fill + d_copy_left + 16) * 2, (size_t)b.fill * 2);
    if (!ok) {
        if (!d_failed) std::fprintf(stderr, "ieee80211 %s: out of host memory for staging, dropping packets\n", d_label.c_str());
        d_failed = true;
        d_copy_left = 0;
        pk.nsamp = 0;
    }
    float* c = b.chan.data() + np * 128;
    for (int i = 0; i < 64; i++) {
        c[2 * i] = i < (int)chan.size() ? chan[i].real() : 0.0f;
        c[2 * i + 1] = i < (int)chan.size() ? chan[i].imag() : 0.0f;
    }
    if (b.ncomplete == 0) b.oldest = std::chrono::steady_clock::now();
    b.pkts.push_back(pk);
    if (d_copy_left == 0) b.ncomplete++;
    return true;
}

int demod_core::ingest(int base, int navail, gr_vector_const_void_star& input_items)
{
    /* items [base, navail) of this call's input window */
    batch_set& b = d_sets[d_cur];
    int consumed = base;
    while (consumed < navail && b.ncomplete < d_env.max_batch) {
        if (d_left == 0) {
            /* packet boundary: the reference waits here until the first item carries tags (demod_impl.cc:70-108) */
            if (!start_packet(d_blk->nitems_read(0) + (uint64_t)consumed)) break;
        }
        const long n = std::min<long>(navail - consumed, d_left);
        const long ncopy = std::min(n, d_copy_left);
        if (ncopy > 0) {
            for (int a = 0; a < d_nant; a++) {
                const float* src = static_cast<const float*>(input_items[a]) + (size_t)consumed * 2;
                stream_copy(b.iq[a].data() + (size_t)b.fill * 2, src, (size_t)ncopy * 8);
            }
            b.fill += (uint64_t)ncopy;
            d_copy_left -= ncopy;
            if (d_copy_left == 0) b.ncomplete++;
        }
        d_left -= n;
        consumed += (int)n;
    }
    return consumed - base;
}

/* The whole packets of the set being filled go to the GPU; returns without waiting for them. */
void demod_core::submit_current()
{
    const double t_in = d_trace ? trace_now() : 0.0;
    struct at_exit {
        demod_core* c;
        double t;
        ~at_exit()
        {
            if (c->d_trace) c->d_t_submit += trace_now() - t;
        }
    } guard{ this, t_in };
    d_n_batches++;
    d_n_pkts += (uint64_t)d_sets[d_cur].ncomplete;
    const int idx = d_cur;
    batch_set& b = d_sets[idx];
    const int npkt = b.ncomplete;
    const bool partial = (int)b.pkts.size() > npkt; /* the packet still arriving moves to the next set */
    b.npkt = npkt;
    b.desc.assign((size_t)npkt, w80211_desc());
    b.emit.clear();
    b.emit_pos = 0;
    b.ndp_chan.clear();
    b.lease = 0;
    b.slot = -1;
    int mode = d_env.handoff;
    if (mode == HANDOFF_AUTO) mode = w80211_lease_consumers(0) > 0 ? HANDOFF_DEVICE : HANDOFF_HOST;
    b.items_on_host = mode != HANDOFF_DEVICE;
    b.want_lease = mode != HANDOFF_HOST;
    bool sent = false;
    if (!d_failed && npkt > 0) {
        /* a slot: at most NSLOTS batches are on the GPU; the oldest is collected (its results wait in its set) */
        int slot = -1;
        for (;;) {
            for (int s = 0; s < NSLOTS; s++)
                if (!d_slot_busy[s]) slot = s;
            if (slot >= 0) break;
            bool any = false;
            for (int q : d_queue)
                if (d_sets[q].state == SET_INFLIGHT) {
                    collect_set(q, true);
                    any = true;
                    break;
                }
            if (!any) break;
        }
        const uint64_t nsamples = partial ? b.pkts.back().iq_offset : b.fill;
        /* at most 416 (one antenna) or 832 (two) LLRs per 72-sample symbol, 20 floats of padding per packet */
        const size_t cap = (size_t)(nsamples * (d_nant == 2 ? 12 : 6)) + (size_t)npkt * 32 + 64;
        if (slot < 0) {
            std::fprintf(stderr, "ieee80211 %s: no free GPU slot, batch dropped\n", d_label.c_str());
        } else if (b.items_on_host && !b.llr.reserve(cap, 0)) {
            std::fprintf(stderr, "ieee80211 %s: out of host memory for %zu LLRs, batch dropped\n", d_label.c_str(), cap);
        } else {
            std::memset(&b.in, 0, sizeof(b.in));
            b.in.npkt = npkt;
            b.in.nant = d_nant;
            b.in.mem = W80211_MEM_HOST;
            for (int a = 0; a < d_nant; a++) b.in.iq[a] = b.iq[a].data();
            b.in.nsamples = nsamples;
            b.in.pkts = b.pkts.data();
            b.in.chan = b.chan.data();
            b.in.mu_pos = d_mupos;
            std::memset(&b.out, 0, sizeof(b.out));
            b.out.mem = W80211_MEM_HOST;
            b.out.desc = b.desc.data();
            if (b.items_on_host) {
                b.out.want_llr = 1;
                b.out.llr = b.llr.data();
                b.out.llr_capacity = b.llr.capacity();
            }
            if (w80211_demod_submit(d_ctx, slot, &b.in, &b.out) == W80211_OK) {
                b.slot = slot;
                d_slot_busy[slot] = true;
                sent = true;
            } else {
                report("w80211_demod_submit");
            }
        }
    }
    if (!sent)
        for (auto& d : b.desc) d.status = W80211_ST_BAD_INPUT;
    b.state = sent ? SET_INFLIGHT : SET_READY;
    d_queue.push_back(idx);
    d_cur = -1;
    if (partial) d_partial_in = idx;
    take_free_set();
}

/* Results of a submitted set: descriptors (and LLRs) are on the host afterwards, the emit list is made. */
bool demod_core::collect_set(int idx, bool wait)
{
    batch_set& b = d_sets[idx];
    if (b.state == SET_READY) return true;
    if (!wait && w80211_rx_poll(d_ctx, b.slot) == 0) return false;
    const double t_w = d_trace ? trace_now() : 0.0;
    int r = w80211_rx_collect(d_ctx, b.slot);
    if (d_trace) d_t_wait += trace_now() - t_w;
    if (r != W80211_OK) {
        report("w80211_demod_submit / w80211_rx_collect");
        for (auto& d : b.desc) d.status = W80211_ST_BAD_INPUT;
    } else if (b.want_lease && b.out.n_llr > 0) {
        if (w80211_llr_export(d_ctx, b.slot, &b.lease) != W80211_OK) {
            b.lease = 0;
            if (!b.items_on_host) {
                report("w80211_llr_export");
                for (auto& d : b.desc) d.status = W80211_ST_BAD_INPUT;
            }
        }
    }
    d_slot_busy[b.slot] = false;
    for (int p = 0; p < b.npkt; p++) {
        /* packets that fail the reference's gates go to its CLEAN state without output (demod_impl.cc:159,186) */
        if (b.desc[p].status == W80211_ST_OK && b.desc[p].total > 0) {
            b.emit.push_back(emit_item{ p, 0, -1 });
        } else if (b.desc[p].status == W80211_ST_NDP && d_nant == 1 && b.desc[p].nss == 2 && b.pkts[p].nsamp >= 392) {
            /* two-stream VHT NDP on the one-antenna block: the 2 x 64 raw samples of the two VHT-LTFs travel
             * to decode in the tag "mu2x1chan" with 1024 placeholder items (demod_impl.cc:238-271,396-398) */
            const std::complex<float>* x = reinterpret_cast<const std::complex<float>*>(b.iq[0].data()) + b.pkts[p].iq_offset;
            std::vector<std::complex<float>> h(128);
            std::copy(x + 248, x + 312, h.begin());
            std::copy(x + 328, x + 392, h.begin() + 64);
            b.ndp_chan.push_back(h);
            b.desc[p].total = 1024;
            b.emit.push_back(emit_item{ p, 0, (int)b.ndp_chan.size() - 1 });
        }
    }
    b.state = SET_READY;
    return true;
}

/* The tags of demod_impl.cc:224-253 / demod2_impl.cc:229-250 on the packet's first LLR. */
void demod_core::write_tags(const batch_set& b, const w80211_desc& d, uint64_t abs_offset, int ndp)
{
    const pmt::pmt_t src = d_blk->alias_pmt();
    auto put_f = [&](const char* k, float v) { d_blk->add_item_tag(0, abs_offset, pmt::mp(k), pmt::from_float(v), src); };
    auto put_l = [&](const char* k, long v) { d_blk->add_item_tag(0, abs_offset, pmt::mp(k), pmt::from_long(v), src); };
    put_f("cfo", d.cfo);
    put_f("snr", d.snr);
    put_f("rssi", d.rssi);
    if (d.format == W80211_F_VHT) {
        put_f("sssnr0", d.sssnr0);
        if (d_nant == 2 && d.nss != 1) put_f("sssnr1", d.sssnr1);
    }
    put_l("format", d.format);
    put_l("mcs", d.mcs);
    put_l("len", d.len);
    put_l("cr", d.cr);
    put_l("ampdu", d.ampdu);
    put_l("trellis", d.trellis);
    if (ndp >= 0) {
        const auto& h = b.ndp_chan[(size_t)ndp];
        d_blk->add_item_tag(0, abs_offset, pmt::mp("mu2x1chan"), pmt::init_c32vector(h.size(), h), src);
    } else if (b.lease) {
        /* private to this repo's blocks: where the packet's LLRs are in HBM (demod_core.h, "LLR hand-off") */
        put_l("w80211_llr", (long)b.lease);
        put_l("w80211_off", (long)d.llr_off);
        if (!b.items_on_host) put_l("w80211_nohost", 1);
    }
    put_l("total", d.total);
}

int demod_core::emit(batch_set& b, float* out, int room, int already)
{
    /* `already`: items written earlier in this general_work() call, before `out` */
    int produced = 0;
    while (b.emit_pos < b.emit.size() && produced < room) {
        emit_item& e = b.emit[b.emit_pos];
        const w80211_desc& d = b.desc[(size_t)e.pkt];
        if (e.written == 0) write_tags(b, d, d_blk->nitems_written(0) + (uint64_t)(already + produced), e.ndp);
        const size_t n = std::min<size_t>((size_t)(room - produced), (size_t)d.total - e.written);
        if (e.ndp >= 0)
            std::memset(out + produced, 0, n * sizeof(float)); /* the reference leaves these items unwritten */
        else if (b.items_on_host)
            std::memcpy(out + produced, b.llr.data() + d.llr_off + e.written, n * sizeof(float));
        /* else: the LLRs stay in HBM (lease tags); the items are produced, not written */
        e.written += n;
        produced += (int)n;
        if (e.written == (size_t)d.total) b.emit_pos++;
    }
    return produced;
}

int demod_core::work(int noutput_items,
                     gr_vector_int& ninput_items,
                     gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items)
{
    float* out = static_cast<float*>(output_items[0]);
    int produced = 0, consumed = 0;
    d_n_work++;
    /* finished batches, oldest first */
    auto drain = [&](bool wait_first) {
        bool wait = wait_first;
        while (!d_queue.empty()) {
            const int idx = d_queue.front();
            batch_set& b = d_sets[idx];
            if (!collect_set(idx, wait)) break;
            wait = false;
            const double t_e = d_trace ? trace_now() : 0.0;
            produced += emit(b, out + produced, noutput_items - produced, produced);
            if (d_trace) d_t_emit += trace_now() - t_e;
            if (b.emit_pos < b.emit.size()) break; /* the output is full */
            d_queue.pop_front();
            if (d_partial_in == idx) {
                /* it still holds the samples of the packet being received (no other set was free): it carries on as the
                 * set being filled */
                w80211_pkt_in last = b.pkts.back();
                const uint64_t have = b.fill - last.iq_offset;
                for (int a = 0; a < d_nant; a++)
                    if (have) std::memmove(b.iq[a].data(), b.iq[a].data() + (size_t)last.iq_offset * 2, (size_t)have * 8);
                std::memmove(b.chan.data(), b.chan.data() + (size_t)b.npkt * 128, 128 * sizeof(float));
                last.iq_offset = 0;
                b.pkts.assign(1, last);
                b.fill = have;
                b.ncomplete = 0;
                b.state = SET_FILLING;
                d_partial_in = -1;
                d_cur = idx;
            } else {
                b.state = SET_FREE;
            }
        }
    };
    drain(false);
    take_free_set();
    int navail = ninput_items[0];
    for (int a = 1; a < d_nant; a++) navail = std::min(navail, ninput_items[a]);
    while (d_cur >= 0) {
        batch_set& b = d_sets[d_cur];
        const double t_i = d_trace ? trace_now() : 0.0;
        consumed += ingest(consumed, navail, input_items);
        if (d_trace) d_t_ingest += trace_now() - t_i;
        if (b.ncomplete == 0) break;
        const bool drained = consumed == navail;
        const double waited = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - b.oldest).count();
        if (b.ncomplete >= d_env.max_batch || (drained && b.ncomplete >= d_env.min_batch) || waited >= d_env.max_wait_ms)
            submit_current();
        else
            break;
        if (drained) break;
    }
    if (produced == 0 && consumed == 0 && !d_queue.empty()) {
        /* nothing else to do: wait for the oldest batch */
        drain(true);
        take_free_set();
    }
    d_blk->consume_each(consumed);
    return produced;
}

} // namespace ieee80211
} // namespace gr
