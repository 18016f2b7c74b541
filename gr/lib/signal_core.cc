/* signal_core.cc -- flag gathering, GPU batch call and packet emission shared by signal and signal2.
 * See signal_core.h for the structure; reference behaviour is cited where it is mirrored. */
#include "signal_core.h"

#include <algorithm>
#include <complex>
#include <cstdio>
#include <cstring>

namespace gr {
namespace ieee80211 {

static const long PAD_ITEMS = 320; /* signal_impl.cc:194-201 */

signal_core::signal_core(gr::block* owner, int nant, const char* label)
    : d_blk(owner),
      d_nant(nant),
      d_label(label),
      d_env(rx_env::read()),
      d_ctx(nullptr),
      d_failed(false),
      d_fill(0),
      d_start(0),
      d_need(0),
      d_emit_pos(0),
      d_emit_written(0),
      d_seq(0)
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
                     "no packet will pass\n",
                     d_label.c_str(), r, d_env.device);
    }
}

signal_core::~signal_core()
{
    if (d_ctx) w80211_destroy(d_ctx);
}

void signal_core::forecast(int /*noutput_items*/, gr_vector_int& ninput_items_required)
{
    const int need = d_emit_pos < d_pkts.size() ? 0 : 1;
    for (auto& n : ninput_items_required) n = need;
}

/* Append the window to the staging and note its sync flags with their tags (signal_impl.cc:80-98:
 * keys "rad", "snr", "rssi", each defaulting to 0; a flag without any tag is skipped by the trigger). */
void signal_core::ingest(int n, gr_vector_const_void_star& input_items)
{
    const uint8_t* sync = static_cast<const uint8_t*>(input_items[0]);
    bool ok = true;
    for (int a = 0; a < d_nant; a++) ok = ok && d_raw[a].reserve((size_t)(d_fill + (uint64_t)n + 16) * 2, (size_t)d_fill * 2);
    if (!ok) {
        if (!d_failed) std::fprintf(stderr, "ieee80211 %s: out of host memory for staging, input dropped\n", d_label.c_str());
        d_failed = true;
    }
    if (d_failed) {
        d_fill = d_start = d_need = 0;
        d_syncs.clear();
        return;
    }
    for (int a = 0; a < d_nant; a++)
        std::memcpy(d_raw[a].data() + (size_t)d_fill * 2, input_items[1 + a], (size_t)n * 8);
    const uint64_t first = d_blk->nitems_read(0);
    for (int i = 0; i < n; i++) {
        if (!sync[i]) continue;
        w80211_sync sy;
        std::memset(&sy, 0, sizeof(sy));
        sy.pos = d_fill + (uint64_t)i;
        std::vector<gr::tag_t> tags;
        d_blk->get_tags_in_range(tags, 0, first + (uint64_t)i, first + (uint64_t)i + 1);
        if (!tags.empty()) {
            sy.flags = W80211_SYNC_TAGGED;
            for (const auto& t : tags) {
                const std::string key = pmt::symbol_to_string(t.key);
                if (key == "rad")
                    sy.rad = (float)pmt::to_float(t.value);
                else if (key == "snr")
                    sy.snr = (float)pmt::to_float(t.value);
                else if (key == "rssi")
                    sy.rssi = (float)pmt::to_float(t.value);
            }
        }
        d_syncs.push_back(sy);
    }
    d_fill += (uint64_t)n;
}

void signal_core::run_batch()
{
    d_pkts.clear();
    d_emit_pos = 0;
    d_emit_written = 0;
    const size_t ns = d_syncs.size();
    w80211_signal_in in;
    std::memset(&in, 0, sizeof(in));
    in.nsync = (int32_t)ns;
    in.nant = d_nant;
    in.mem = W80211_MEM_HOST;
    in.final = 0;
    for (int a = 0; a < d_nant; a++) in.iq[a] = d_raw[a].data();
    in.nsamples = d_fill;
    in.syncs = d_syncs.data();
    in.start = std::min(d_start, d_fill);
    w80211_signal_out out;
    std::memset(&out, 0, sizeof(out));
    out.mem = W80211_MEM_HOST;
    bool ok = d_chan.reserve(ns * 128, 0);
    for (int a = 0; a < d_nant; a++) ok = ok && d_out[a].reserve((size_t)(d_fill + 16) * 2, 0);
    d_pkts.resize(ns);
    out.pkts = d_pkts.data();
    out.chan = d_chan.data();
    for (int a = 0; a < d_nant; a++) out.iq[a] = d_out[a].data();
    out.iq_capacity = d_fill;
    int r = ok ? w80211_signal_batch(d_ctx, &in, &out) : W80211_E_NOMEM;
    if (r != W80211_OK) {
        std::fprintf(stderr, "ieee80211 %s: w80211_signal_batch: %s; %zu flags dropped\n", d_label.c_str(),
                     ok ? w80211_last_error(d_ctx) : "out of host memory", ns);
        d_pkts.clear();
        d_syncs.clear();
        d_fill = d_start = d_need = 0;
        return;
    }
    d_pkts.resize((size_t)out.npkt);
    /* drop what the state machine has passed; the unsettled flag (if any) is now at position 0 */
    const uint64_t next = std::min<uint64_t>(out.next, d_fill);
    const uint64_t keep = d_fill - next;
    if (keep && next)
        for (int a = 0; a < d_nant; a++) std::memmove(d_raw[a].data(), d_raw[a].data() + (size_t)next * 2, (size_t)keep * 8);
    d_syncs.erase(d_syncs.begin(), d_syncs.begin() + out.nsync_used);
    for (auto& sy : d_syncs) sy.pos -= next;
    d_fill = keep;
    d_start = 0;
    d_need = out.need > next ? out.need - next : 0;
}

/* The tags of signal_impl.cc:134-152 on the packet's first item (output 0 only, also for signal2). */
void signal_core::write_tags(const w80211_pkt_in& pk, const float* chan, long seq, uint64_t abs_offset)
{
    const pmt::pmt_t src = d_blk->alias_pmt();
    auto put = [&](const char* k, const pmt::pmt_t& v) { d_blk->add_item_tag(0, abs_offset, pmt::mp(k), v, src); };
    std::vector<std::complex<float>> h(64);
    for (int i = 0; i < 64; i++) h[i] = std::complex<float>(chan[2 * i], chan[2 * i + 1]);
    put("cfo", pmt::from_float(pk.cfo));
    put("snr", pmt::from_float(pk.snr));
    put("rssi", pmt::from_float(pk.rssi));
    put("seq", pmt::from_long(seq));
    put("mcs", pmt::from_long(pk.l_mcs));
    put("len", pmt::from_long(pk.l_len));
    put("nsamp", pmt::from_long(pk.nsamp));
    put("chan", pmt::init_c32vector(h.size(), h));
}

int signal_core::emit(gr_vector_void_star& output_items, int room, int already)
{
    int produced = 0;
    while (d_emit_pos < d_pkts.size() && produced < room) {
        const w80211_pkt_in& pk = d_pkts[d_emit_pos];
        if (d_emit_written == 0) {
            d_seq++;
            if (d_seq >= 1000000000) d_seq = 0;
            write_tags(pk, d_chan.data() + d_emit_pos * 128, d_seq, d_blk->nitems_written(0) + (uint64_t)(already + produced));
        }
        const long total = (long)pk.nsamp + PAD_ITEMS;
        const long n = std::min<long>(room - produced, total - d_emit_written);
        const long nsig = std::max(0L, std::min(n, (long)pk.nsamp - d_emit_written));
        for (int a = 0; a < d_nant; a++) {
            float* o = static_cast<float*>(output_items[a]) + (size_t)(already + produced) * 2;
            if (nsig) std::memcpy(o, d_out[a].data() + (size_t)(pk.iq_offset + (uint64_t)d_emit_written) * 2, (size_t)nsig * 8);
            if (n > nsig) std::memset(o + (size_t)nsig * 2, 0, (size_t)(n - nsig) * 8); /* pad items (left unwritten by the reference) */
        }
        d_emit_written += n;
        produced += (int)n;
        if (d_emit_written == total) {
            d_emit_pos++;
            d_emit_written = 0;
        }
    }
    if (d_emit_pos >= d_pkts.size()) {
        d_pkts.clear();
        d_emit_pos = 0;
    }
    return produced;
}

int signal_core::work(int noutput_items,
                      gr_vector_int& ninput_items,
                      gr_vector_const_void_star& input_items,
                      gr_vector_void_star& output_items)
{
    int produced = emit(output_items, noutput_items, 0);
    int consumed = 0;
    if (d_pkts.empty()) {
        int n = ninput_items[0];
        for (int a = 0; a < d_nant; a++) n = std::min(n, ninput_items[1 + a]);
        if (n > 0) {
            ingest(n, input_items);
            consumed = n;
            if (d_syncs.empty()) {
                /* TRIGGER state with no flag in sight: everything is consumed (signal_impl.cc:76-107) */
                d_fill = d_start = d_need = 0;
            } else if (d_fill >= d_need) {
                run_batch();
                produced += emit(output_items, noutput_items - produced, produced);
            }
        }
    }
    d_blk->consume_each(consumed);
    return produced;
}

} // namespace ieee80211
} // namespace gr
