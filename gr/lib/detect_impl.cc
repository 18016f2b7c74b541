/* detect_impl.cc -- see detect_impl.h; the stream protocol of the output is the reference sync block's
 * (lib/sync_impl.cc:61-149). */
#include "detect_impl.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include <gnuradio/io_signature.h>

namespace gr {
namespace ieee80211 {

detect::sptr detect::make() { return gnuradio::make_block_sptr<detect_impl>(); }

detect_impl::detect_impl()
    : gr::block("detect", gr::io_signature::make(1, 1, sizeof(gr_complex)), gr::io_signature::make(1, 1, sizeof(uint8_t))),
      d_env(rx_env::read()),
      d_ctx(nullptr),
      d_failed(false),
      d_chunk(1u << 20),
      d_have(0),
      d_lead(0),
      d_fresh(false),
      d_have_state(false),
      d_examined(0),
      d_emitted(0)
{
    set_tag_propagation_policy(block::TPP_DONT);
    std::memset(&d_state, 0, sizeof(d_state));
    if (const char* c = std::getenv("W80211_DETECT_CHUNK")) d_chunk = (size_t)std::max(512L, std::atol(c));
    w80211_cfg cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.device = d_env.device;
    cfg.nslots = 1;
    int r = w80211_abi_version() == W80211_ABI_VERSION ? w80211_create(&cfg, &d_ctx) : W80211_E_ARG;
    if (r != W80211_OK) {
        d_ctx = nullptr;
        d_failed = true;
        std::fprintf(stderr,
                     "ieee80211 detect: w80211_create failed (%d): no usable CUDA device %d; there is no CPU path, no packet "
                     "will be flagged\n",
                     r, d_env.device);
    }
}

detect_impl::~detect_impl()
{
    if (d_ctx) w80211_destroy(d_ctx);
}

void detect_impl::forecast(int /*noutput_items*/, gr_vector_int& ninput_items_required)
{
    /* decided but unwritten output, or samples waiting for their GPU call, need no new input */
    ninput_items_required[0] = (d_emitted < d_examined || (d_fresh && d_have > d_lead)) ? 0 : 1;
}

/* One GPU call over the buffer: flags of the examined samples are queued, the unexamined tail and 96 samples of history
 * before it stay. */
void detect_impl::process()
{
    d_fresh = false;
    const size_t n = d_have - d_lead;
    if (n == 0) return;
    uint64_t next = n;
    if (!d_failed) {
        d_syncs.resize(n / 100 + 16);
        w80211_detect_out out;
        std::memset(&out, 0, sizeof(out));
        out.syncs = d_syncs.data();
        out.sync_capacity = (int32_t)d_syncs.size();
        int r = w80211_detect_stream(d_ctx, d_x.data(), d_have, d_lead, W80211_MEM_HOST, 0, d_have_state ? &d_state : nullptr, &out);
        if (r != W80211_OK) {
            std::fprintf(stderr, "ieee80211 detect: w80211_detect_stream: %s; %zu samples pass unflagged\n", w80211_last_error(d_ctx), n);
            std::memset(&d_state, 0, sizeof(d_state));
            d_have_state = false;
        } else {
            next = out.next;
            d_state = out.state;
            d_have_state = true;
            for (int k = 0; k < out.nsync; k++) {
                const w80211_sync& s = d_syncs[(size_t)k];
                d_flags.push_back(flag{ d_examined + s.pos, s.rad, s.snr, s.rssi });
            }
        }
    }
    d_examined += next;
    /* keep the unexamined tail with its history */
    const size_t done = d_lead + (size_t)next;           /* samples of d_x before the first unexamined one */
    const size_t keep_hist = std::min<size_t>(done, 96); /* presiso needs 79 */
    const size_t from = done - keep_hist;
    if (from > 0) std::memmove(d_x.data(), d_x.data() + 2 * from, (d_have - from) * 2 * sizeof(float));
    d_have -= from;
    d_lead = keep_hist;
}

int detect_impl::general_work(int noutput_items,
                              gr_vector_int& ninput_items,
                              gr_vector_const_void_star& input_items,
                              gr_vector_void_star& output_items)
{
    uint8_t* out = static_cast<uint8_t*>(output_items[0]);
    const float* in = static_cast<const float*>(input_items[0]);
    const int navail = ninput_items[0];
    int consumed = 0, produced = 0;
    auto emit = [&]() {
        const uint64_t n = std::min<uint64_t>(d_examined - d_emitted, (uint64_t)(noutput_items - produced));
        if (n == 0) return;
        std::memset(out + produced, 0, (size_t)n);
        const pmt::pmt_t src = alias_pmt();
        while (!d_flags.empty() && d_flags.front().pos < d_emitted + n) {
            const flag f = d_flags.front();
            d_flags.pop_front();
            out[produced + (f.pos - d_emitted)] = 0x01; /* sync index is LTF starting index + 16 (sync_impl.cc:121) */
            const uint64_t at = nitems_written(0) + (uint64_t)produced + (f.pos - d_emitted);
            add_item_tag(0, at, pmt::mp("rad"), pmt::from_float(f.rad), src);
            add_item_tag(0, at, pmt::mp("snr"), pmt::from_float(f.snr), src);
            add_item_tag(0, at, pmt::mp("rssi"), pmt::from_float(f.rssi), src);
        }
        d_emitted += n;
        produced += (int)n;
    };
    emit();
    while (consumed < navail) {
        const size_t room = d_lead + d_chunk > d_have ? d_lead + d_chunk - d_have : 0;
        const size_t n = std::min<size_t>(room, (size_t)(navail - consumed));
        if (n > 0) {
            if (!d_x.reserve((d_have + n + 16) * 2, d_have * 2)) {
                if (!d_failed) std::fprintf(stderr, "ieee80211 detect: out of host memory for staging\n");
                d_failed = true;
                break;
            }
            stream_copy(d_x.data() + 2 * d_have, in + 2 * (size_t)consumed, n * 2 * sizeof(float));
            d_have += n;
            consumed += (int)n;
            d_fresh = true;
        }
        if (d_have - d_lead >= d_chunk) {
            const size_t before = d_have - d_lead;
            process();
            if (d_have - d_lead >= before) break; /* nothing could be examined (a trigger waits for samples beyond the chunk) */
        } else if (n == 0) {
            break;
        }
    }
    /* the input has run dry: examine what is there rather than wait for a full chunk */
    if (consumed == navail && d_fresh && d_have > d_lead) process();
    emit();
    consume_each(consumed);
    return produced;
}

} // namespace ieee80211
} // namespace gr
