/* demod_core.cc -- packet gathering, GPU batch call and LLR emission shared by demod and demod2.
 * See demod_core.h for the structure; reference behaviour is cited where it is mirrored. */
#include "demod_core.h"

#include <algorithm>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <cstring>

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
      d_fill(0),
      d_ncomplete(0),
      d_left(0),
      d_copy_left(0),
      d_emit_pos(0)
{
    w80211_cfg cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.device = d_env.device;
    cfg.max_packets = d_env.max_batch;
    cfg.nslots = 1;
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
}

demod_core::~demod_core()
{
    if (d_ctx) w80211_destroy(d_ctx);
}

void demod_core::report(const char* what)
{
    std::fprintf(stderr, "ieee80211 %s: %s: %s; batch dropped\n", d_label.c_str(), what, w80211_last_error(d_ctx));
}

void demod_core::forecast(int /*noutput_items*/, gr_vector_int& ninput_items_required)
{
    /* queued LLRs or whole staged packets can be turned into output without new input */
    const int need = flush_pending() ? 0 : 1;
    for (auto& n : ninput_items_required) n = need;
}

/* Tags on the first item of a packet, keys and defaults as at demod_impl.cc:70-89. */
bool demod_core::start_packet(uint64_t abs_offset)
{
    std::vector<gr::tag_t> tags;
    d_blk->get_tags_in_range(tags, 0, abs_offset, abs_offset + 1);
    if (tags.empty()) return false;
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
    d_left = nsamp + 320; /* demod_impl.cc:92 */
    if (d_left <= 0) return false; /* the reference never leaves this tag either */
    d_copy_left = std::max(nsamp, 0L);
    pk.nsamp = (int32_t)d_copy_left;
    pk.iq_offset = d_fill;
    const size_t np = d_pkts.size();
    bool ok = d_chan.reserve((np + 1) * 128, np * 128);
    for (int a = 0; a < d_nant; a++) ok = ok && d_iq[a].reserve((size_t)(d_fill + d_copy_left + 16) * 2, (size_t)d_fill * 2);
    if (!ok) {
        if (!d_failed) std::fprintf(stderr, "ieee80211 %s: out of host memory for staging, dropping packets\n", d_label.c_str());
        d_failed = true;
        d_copy_left = 0;
        pk.nsamp = 0;
    }
    float* c = d_chan.data() + np * 128;
    for (int i = 0; i < 64; i++) {
        c[2 * i] = i < (int)chan.size() ? chan[i].real() : 0.0f;
        c[2 * i + 1] = i < (int)chan.size() ? chan[i].imag() : 0.0f;
    }
    if (d_ncomplete == 0) d_oldest = std::chrono::steady_clock::now();
    d_pkts.push_back(pk);
    if (d_copy_left == 0) d_ncomplete++;
    return true;
}

int demod_core::ingest(int navail, gr_vector_const_void_star& input_items)
{
    int consumed = 0;
    while (consumed < navail && d_ncomplete < d_env.max_batch) {
        if (d_left == 0) {
            /* packet boundary: the reference waits here until the first item carries tags (demod_impl.cc:70-108) */
            if (!start_packet(d_blk->nitems_read(0) + (uint64_t)consumed)) break;
        }
        const long n = std::min<long>(navail - consumed, d_left);
        const long ncopy = std::min(n, d_copy_left);
        if (ncopy > 0) {
            for (int a = 0; a < d_nant; a++) {
                const float* src = static_cast<const float*>(input_items[a]) + (size_t)consumed * 2;
                std::memcpy(d_iq[a].data() + (size_t)d_fill * 2, src, (size_t)ncopy * 8);
            }
            d_fill += (uint64_t)ncopy;
            d_copy_left -= ncopy;
            if (d_copy_left == 0) d_ncomplete++;
        }
        d_left -= n;
        consumed += (int)n;
    }
    return consumed;
}

void demod_core::run_batch()
{
    const int npkt = d_ncomplete;
    const bool partial = (int)d_pkts.size() > npkt; /* the packet still arriving stays staged */
    d_desc.assign((size_t)npkt, w80211_desc());
    d_emit.clear();
    d_emit_pos = 0;
    d_ndp_chan.clear();
    if (!d_failed && npkt > 0) {
        uint64_t nsamp_tot = 0;
        for (int p = 0; p < npkt; p++) nsamp_tot += (uint64_t)d_pkts[p].nsamp;
        /* at most 416 (one antenna) or 832 (two) LLRs per 72-sample symbol, 20 floats of padding per packet */
        const size_t cap = (size_t)(nsamp_tot * (d_nant == 2 ? 12 : 6)) + (size_t)npkt * 32 + 64;
        if (!d_llr.reserve(cap, 0)) {
            std::fprintf(stderr, "ieee80211 %s: out of host memory for %zu LLRs, batch dropped\n", d_label.c_str(), cap);
        } else {
            w80211_batch_in in;
            std::memset(&in, 0, sizeof(in));
            in.npkt = npkt;
            in.nant = d_nant;
            in.mem = W80211_MEM_HOST;
            for (int a = 0; a < d_nant; a++) in.iq[a] = d_iq[a].data();
            in.nsamples = d_fill;
            in.pkts = d_pkts.data();
            in.chan = d_chan.data();
            in.mu_pos = d_mupos;
            w80211_batch_out out;
            std::memset(&out, 0, sizeof(out));
            out.mem = W80211_MEM_HOST;
            out.want_llr = 1;
            out.desc = d_desc.data();
            out.llr = d_llr.data();
            out.llr_capacity = d_llr.capacity();
            int r = w80211_demod_batch(d_ctx, &in, &out);
            if (r != W80211_OK) {
                report("w80211_demod_batch");
                for (auto& d : d_desc) d.status = W80211_ST_BAD_INPUT;
            }
            for (int p = 0; p < npkt; p++) {
                /* packets that fail the reference's gates go to its CLEAN state without output (demod_impl.cc:159,186) */
                if (d_desc[p].status == W80211_ST_OK && d_desc[p].total > 0) {
                    d_emit.push_back(emit_item{ p, 0, -1 });
                } else if (d_desc[p].status == W80211_ST_NDP && d_nant == 1 && d_desc[p].nss == 2 && d_pkts[p].nsamp >= 392) {
                    /* two-stream VHT NDP on the one-antenna block: the 2 x 64 raw samples of the two VHT-LTFs travel
                     * to decode in the tag "mu2x1chan" with 1024 placeholder items (demod_impl.cc:238-271,396-398) */
                    const std::complex<float>* x = reinterpret_cast<const std::complex<float>*>(d_iq[0].data()) + d_pkts[p].iq_offset;
                    std::vector<std::complex<float>> h(128);
                    std::copy(x + 248, x + 312, h.begin());
                    std::copy(x + 328, x + 392, h.begin() + 64);
                    d_ndp_chan.push_back(h);
                    d_desc[p].total = 1024;
                    d_emit.push_back(emit_item{ p, 0, (int)d_ndp_chan.size() - 1 });
                }
            }
        }
    }
    /* keep the packet still arriving */
    if (partial) {
        w80211_pkt_in last = d_pkts.back();
        const uint64_t have = d_fill - last.iq_offset;
        for (int a = 0; a < d_nant; a++)
            if (have) std::memmove(d_iq[a].data(), d_iq[a].data() + (size_t)last.iq_offset * 2, (size_t)have * 8);
        std::memmove(d_chan.data(), d_chan.data() + (size_t)npkt * 128, 128 * sizeof(float));
        last.iq_offset = 0;
        d_pkts.assign(1, last);
        d_fill = have;
    } else {
        d_pkts.clear();
        d_fill = 0;
    }
    d_ncomplete = 0;
}

/* The tags of demod_impl.cc:224-253 / demod2_impl.cc:229-250 on the packet's first LLR. */
void demod_core::write_tags(const w80211_desc& d, uint64_t abs_offset, int ndp)
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
        const auto& h = d_ndp_chan[(size_t)ndp];
        d_blk->add_item_tag(0, abs_offset, pmt::mp("mu2x1chan"), pmt::init_c32vector(h.size(), h), src);
    }
    put_l("total", d.total);
}

int demod_core::emit(float* out, int room, int already)
{
    /* `already`: items written earlier in this general_work() call, before `out` */
    int produced = 0;
    while (d_emit_pos < d_emit.size() && produced < room) {
        emit_item& e = d_emit[d_emit_pos];
        const w80211_desc& d = d_desc[(size_t)e.pkt];
        if (e.written == 0) write_tags(d, d_blk->nitems_written(0) + (uint64_t)(already + produced), e.ndp);
        const size_t n = std::min<size_t>((size_t)(room - produced), (size_t)d.total - e.written);
        if (e.ndp >= 0)
            std::memset(out + produced, 0, n * sizeof(float)); /* the reference leaves these items unwritten */
        else
            std::memcpy(out + produced, d_llr.data() + d.llr_off + e.written, n * sizeof(float));
        e.written += n;
        produced += (int)n;
        if (e.written == (size_t)d.total) d_emit_pos++;
    }
    if (d_emit_pos >= d_emit.size()) {
        d_emit.clear();
        d_emit_pos = 0;
    }
    return produced;
}

int demod_core::work(int noutput_items,
                     gr_vector_int& ninput_items,
                     gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items)
{
    float* out = static_cast<float*>(output_items[0]);
    int produced = emit(out, noutput_items, 0);
    int consumed = 0;
    if (d_emit.empty()) {
        int navail = ninput_items[0];
        for (int a = 1; a < d_nant; a++) navail = std::min(navail, ninput_items[a]);
        consumed = ingest(navail, input_items);
        if (d_ncomplete > 0) {
            const bool drained = consumed == navail;
            const double waited =
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - d_oldest).count();
            if (d_ncomplete >= d_env.max_batch || (drained && d_ncomplete >= d_env.min_batch) || waited >= d_env.max_wait_ms) {
                run_batch();
                produced += emit(out + produced, noutput_items - produced, produced);
            }
        }
    }
    d_blk->consume_each(consumed);
    return produced;
}

} // namespace ieee80211
} // namespace gr
