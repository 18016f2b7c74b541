/* decode_impl.cc -- decoder block: stream and message protocol of the reference
 * (lib/decode_impl.cc:33-162), batches decoded on the GPU through w80211_decode_submit() without waiting. */
#include "decode_impl.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>

#include <gnuradio/io_signature.h>

namespace gr {
namespace ieee80211 {

decode::sptr decode::make(bool ifdebug) { return gnuradio::make_block_sptr<decode_impl>(ifdebug); }

decode_impl::decode_impl(bool ifdebug)
    : gr::block("decode", gr::io_signature::make(1, 1, sizeof(float)), gr::io_signature::make(0, 0, 0)),
      d_debug(ifdebug),
      d_env(rx_env::read()),
      d_ctx(nullptr),
      d_failed(false),
      d_cur(-1),
      d_partial_in(-1),
      d_left(0),
      d_copy_left(0),
      d_lost_leases(0),
      d_trace(std::getenv("W80211_TRACE") != nullptr),
      d_t_ingest(0),
      d_t_submit(0),
      d_t_wait(0),
      d_t_publish(0),
      d_n_work(0),
      d_n_batches(0),
      d_n_pkts(0),
      d_n_leased(0),
      d_nPktCorrect(0)
{
    message_port_register_out(pmt::mp("out"));
    std::memset(d_legacyMcsCount, 0, sizeof(d_legacyMcsCount));
    std::memset(d_htMcsCount, 0, sizeof(d_htMcsCount));
    std::memset(d_vhtMcsCount, 0, sizeof(d_vhtMcsCount));
    for (bool& b : d_slot_busy) b = false;
    set_tag_propagation_policy(block::TPP_DONT);
    w80211_cfg cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.device = d_env.device;
    cfg.max_packets = d_env.max_batch;
    cfg.nslots = NSLOTS;
    int r = w80211_abi_version() == W80211_ABI_VERSION ? w80211_create(&cfg, &d_ctx) : W80211_E_ARG;
    if (w80211_abi_version() != W80211_ABI_VERSION)
        std::fprintf(stderr, "ieee80211 %s: libw80211rx.so has ABI version %d, this block was built for %d\n", "decode",
                     w80211_abi_version(), W80211_ABI_VERSION);
    if (r != W80211_OK) {
        d_ctx = nullptr;
        d_failed = true;
        std::fprintf(stderr,
                     "ieee80211 decode: w80211_create failed (%d): no usable CUDA device %d; there is no CPU path, "
                     "packets will be dropped\n",
                     r, d_env.device);
    } else {
        w80211_lease_consumers(+1); /* demod blocks of this process may leave their LLRs in HBM from now on */
    }
    take_free_set();
}

decode_impl::~decode_impl()
{
    if (d_trace)
        std::fprintf(stderr, "ieee80211 decode trace: %llu general_work calls, %llu batches, %llu packets (%llu from LLR leases, %llu lost "
                             "leases); ingest %.1f ms, submit %.1f ms, waiting for the GPU %.1f ms, publish %.1f ms\n",
                     (unsigned long long)d_n_work, (unsigned long long)d_n_batches, (unsigned long long)d_n_pkts, (unsigned long long)d_n_leased,
                     (unsigned long long)d_lost_leases, 1e3 * d_t_ingest, 1e3 * d_t_submit, 1e3 * d_t_wait, 1e3 * d_t_publish);
    if (d_ctx) {
        w80211_lease_consumers(-1);
        for (int idx : d_queue)
            if (d_sets[idx].state == SET_INFLIGHT) w80211_rx_collect(d_ctx, d_sets[idx].slot);
        w80211_destroy(d_ctx);
    }
}

void decode_impl::forecast(int /*noutput_items*/, gr_vector_int& ninput_items_required)
{
    const bool pending = !d_queue.empty() || (d_cur >= 0 && d_sets[d_cur].ncomplete > 0);
    ninput_items_required[0] = pending ? 0 : 1;
}

bool decode_impl::take_free_set()
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
    b.desc.clear();
    b.fill = 0;
    b.ncomplete = 0;
    b.lease = 0;
    if (d_partial_in >= 0) {
        /* the packet that was still arriving when that set was submitted: its LLRs so far move over */
        batch_set& src = d_sets[d_partial_in];
        w80211_desc last = src.desc.back();
        const uint64_t have = src.fill - (uint64_t)last.llr_off;
        if (b.llr.reserve((size_t)last.total + 64, 0)) {
            if (have) std::memcpy(b.llr.data(), src.llr.data() + last.llr_off, (size_t)have * sizeof(float));
        } else {
            if (!d_failed) std::fprintf(stderr, "ieee80211 decode: out of host memory for staging, dropping packets\n");
            d_failed = true;
        }
        last.llr_off = 0;
        b.desc.push_back(last);
        b.fill = have;
        d_partial_in = -1;
    }
    d_cur = idx;
    return true;
}

/* Tags on the first LLR of a packet, keys and defaults as at decode_impl.cc:70-89, gates as at :93-121. */
int decode_impl::start_packet(uint64_t abs_offset)
{
    std::vector<gr::tag_t> tags;
    get_tags_in_range(tags, 0, abs_offset, abs_offset + 1);
    if (tags.empty()) return START_NONE;
    w80211_desc d;
    std::memset(&d, 0, sizeof(d));
    long format = 99999, len = 99999, total = 99999, cr = 99999, mcs = 99999, ampdu = 99999, trellis = 99999;
    long lease_tag = 0, lease_off = -1, nohost = 0;
    pmt::pmt_t ndp_chan = pmt::PMT_NIL;
    for (const auto& t : tags) {
        const std::string key = pmt::symbol_to_string(t.key);
        if (key == "format") format = pmt::to_long(t.value);
        else if (key == "len") len = pmt::to_long(t.value);
        else if (key == "total") total = pmt::to_long(t.value);
        else if (key == "cr") cr = pmt::to_long(t.value);
        else if (key == "mcs") mcs = pmt::to_long(t.value);
        else if (key == "ampdu") ampdu = pmt::to_long(t.value);
        else if (key == "trellis") trellis = pmt::to_long(t.value);
        else if (key == "cfo") d.cfo = (float)pmt::to_float(t.value);
        else if (key == "snr") d.snr = (float)pmt::to_float(t.value);
        else if (key == "rssi") d.rssi = (float)pmt::to_float(t.value);
        else if (key == "sssnr0") d.sssnr0 = (float)pmt::to_float(t.value);
        else if (key == "sssnr1") d.sssnr1 = (float)pmt::to_float(t.value);
        else if (key == "mu2x1chan") ndp_chan = t.value;
        else if (key == "w80211_llr") lease_tag = pmt::to_long(t.value);
        else if (key == "w80211_off") lease_off = pmt::to_long(t.value);
        else if (key == "w80211_nohost") nohost = pmt::to_long(t.value);
    }
    if (total <= 0) return START_NONE; /* malformed stream: nothing to skip over */
    if (len > W80211_B_MAX || trellis > W80211_T_MAX) { /* skipped by the reference, decode_impl.cc:93-97 */
        d_left = total;
        d_copy_left = 0;
        return START_OK;
    }
    if (trellis == 0) {
        /* VHT NDP: forward the 2x1 channel of the tag as [20, len lo, len hi, 256 floats] (decode_impl.cc:100-121).
         * Every earlier packet is decoded and published first so the publication order stays the arrival order. */
        if (d_sets[d_cur].ncomplete > 0) submit_current();
        drain_all();
        if (!take_free_set()) return START_LATER;
        d_left = total;
        d_copy_left = 0;
        std::vector<std::complex<float>> ch;
        if (!pmt::is_null(ndp_chan)) ch = pmt::c32vector_elements(ndp_chan);
        const int nbytes = 256 * (int)sizeof(float);
        d_msg.assign((size_t)nbytes + 3, 0);
        d_msg[0] = 20; /* C8P_F_VHT_CHAN, lib/cloud80211phy.h */
        d_msg[1] = (unsigned char)(nbytes % 256);
        d_msg[2] = (unsigned char)(nbytes / 256);
        for (int i = 0; i < 128 && i < (int)ch.size(); i++) {
            const float re = ch[(size_t)i].real(), im = ch[(size_t)i].imag();
            std::memcpy(&d_msg[3 + 8 * (size_t)i], &re, 4);
            std::memcpy(&d_msg[3 + 8 * (size_t)i + 4], &im, 4);
        }
        pmt::pmt_t meta = pmt::dict_add(pmt::make_dict(), pmt::mp("len"), pmt::from_long(nbytes + 3));
        message_port_pub(pmt::mp("out"), pmt::cons(meta, pmt::make_blob(d_msg.data(), d_msg.size())));
        return START_OK;
    }
    if (trellis < 0 || cr < 0 || cr > 3 || len < 0) { /* cannot be decoded; skip its LLRs */
        d_left = total;
        d_copy_left = 0;
        return START_OK;
    }
    /* where this packet's LLRs are: a lease of this repo's demod (still valid, on this GPU), or the stream items */
    uint64_t src = 0;
    if (lease_tag > 0 && lease_off >= 0 && (lease_off & 3) == 0 && !d_failed) {
        int32_t dev = -1;
        uint64_t nf = 0;
        if (w80211_llr_query((uint64_t)lease_tag, &dev, &nf) == W80211_OK && dev == d_env.device && (uint64_t)lease_off + (uint64_t)total <= nf)
            src = (uint64_t)lease_tag;
    }
    if (!src && nohost) {
        /* the items were never written and the lease is gone (this block fell more than W80211_LEASES_KEPT batches behind) */
        if (d_lost_leases++ == 0)
            std::fprintf(stderr, "ieee80211 decode: a packet's LLRs were left on the GPU (W80211_LLR_HANDOFF=device) and their lease has "
                                 "been retired; packet dropped (further ones are counted silently)\n");
        d_left = total;
        d_copy_left = 0;
        return START_OK;
    }
    if (!d_sets[d_cur].desc.empty() && d_sets[d_cur].lease != src) {
        /* one source per batch */
        submit_current();
        if (d_cur < 0) return START_LATER;
    }
    batch_set& b = d_sets[d_cur];
    b.lease = src;
    d.status = W80211_ST_OK;
    d.format = (int32_t)format;
    d.mcs = (int32_t)mcs;
    d.len = (int32_t)len;
    d.cr = (int32_t)cr;
    d.ampdu = (int32_t)ampdu;
    d.trellis = (int32_t)trellis;
    d.total = (int32_t)total;
    d_left = total;
    if (b.ncomplete == 0) b.oldest = std::chrono::steady_clock::now();
    if (src) {
        d.llr_off = (int64_t)lease_off;
        d_copy_left = 0;
        b.desc.push_back(d);
        b.ncomplete++;
        return START_OK;
    }
    d.llr_off = (int64_t)((b.fill + 3) & ~(uint64_t)3); /* 16-byte aligned: decoded from where it is staged */
    if (!b.llr.reserve((size_t)d.llr_off + (size_t)total + 64, (size_t)b.fill)) {
        if (!d_failed) std::fprintf(stderr, "ieee80211 decode: out of host memory for staging, dropping packets\n");
        d_failed = true;
        d_copy_left = 0;
        return START_OK;
    }
    b.fill = (uint64_t)d.llr_off;
    d_copy_left = total;
    b.desc.push_back(d);
    return START_OK;
}

void decode_impl::debug_line(const w80211_desc& d, bool crc_ok)
{
    /* same text as the reference prints (decode_impl.cc:376-411, 455-477, 481-509) */
    std::string s("ieee80211 decode, ");
    const uint64_t* cnt;
    int ncnt;
    if (d.format == W80211_F_VHT) {
        s += "vht";
        cnt = d_vhtMcsCount;
        ncnt = 10;
    } else if (d.format == W80211_F_L) {
        s += "legacy";
        cnt = d_legacyMcsCount;
        ncnt = 8;
    } else {
        s += "ht";
        cnt = d_htMcsCount;
        ncnt = 8;
    }
    s += crc_ok ? " crc32 correct, total:" : " crc32 wrong, total:";
    s += std::to_string(d_nPktCorrect);
    for (int i = 0; i < ncnt; i++) s += "," + std::to_string(i) + ":" + std::to_string(cnt[i]);
    s += ",cfo:" + std::to_string(d.cfo) + ",snr:" + std::to_string(d.snr) + ",rssi:" + std::to_string(d.rssi);
    if (d.format == W80211_F_VHT) s += ",sssnr0:" + std::to_string(d.sssnr0) + ",sssnr1:" + std::to_string(d.sssnr1);
    std::cout << s << std::endl;
}

void decode_impl::publish(const w80211_desc& d, const uint8_t* mpdu, uint32_t len, bool crc_ok)
{
    if (d_debug) {
        if (crc_ok) {
            d_nPktCorrect++;
            if (d.format == W80211_F_VHT) {
                if (d.mcs >= 0 && d.mcs < 10) d_vhtMcsCount[d.mcs]++;
            } else if (d.format == W80211_F_L) {
                if (d.mcs >= 0 && d.mcs < 8) d_legacyMcsCount[d.mcs]++;
            } else if (d.mcs >= 0) {
                d_htMcsCount[d.mcs % 8]++;
            }
        }
        debug_line(d, crc_ok);
    }
    if (!crc_ok) return; /* only FCS-valid frames are published (decode_impl.cc:374,453) */
    d_msg.resize((size_t)len + 4);
    d_msg[0] = (unsigned char)d.format;
    d_msg[1] = (unsigned char)(len % 256);
    d_msg[2] = (unsigned char)(len / 256);
    std::memcpy(&d_msg[3], mpdu, len);
    d_msg[3 + (size_t)len] = (unsigned char)d.mcs;
    pmt::pmt_t meta = pmt::dict_add(pmt::make_dict(), pmt::mp("len"), pmt::from_long((long)len + 4));
    message_port_pub(pmt::mp("out"), pmt::cons(meta, pmt::make_blob(d_msg.data(), d_msg.size())));
}

/* The whole packets of the set being filled go to the GPU; returns without waiting for them. */
void decode_impl::submit_current()
{
    const double t_in = d_trace ? trace_now() : 0.0;
    struct at_exit {
        decode_impl* c;
        double t;
        ~at_exit()
        {
            if (c->d_trace) c->d_t_submit += trace_now() - t;
        }
    } guard{ this, t_in };
    d_n_batches++;
    d_n_pkts += (uint64_t)d_sets[d_cur].ncomplete;
    if (d_sets[d_cur].lease) d_n_leased += (uint64_t)d_sets[d_cur].ncomplete;
    const int idx = d_cur;
    batch_set& b = d_sets[idx];
    const int npkt = b.ncomplete;
    const bool partial = (int)b.desc.size() > npkt; /* the packet still arriving moves to the next set */
    b.npkt = npkt;
    b.ok = false;
    b.slot = -1;
    bool sent = false;
    if (!d_failed && npkt > 0) {
        int slot = -1;
        for (;;) {
            for (int s = 0; s < NSLOTS; s++)
                if (!d_slot_busy[s]) slot = s;
            if (slot >= 0 || d_queue.empty()) break;
            finish_front(true); /* both slots are busy: the oldest batch is collected and published first */
        }
        size_t bytes_cap = 64, mpdu_cap = 64;
        for (int p = 0; p < npkt; p++) {
            bytes_cap += (size_t)((b.desc[(size_t)p].trellis + 7) / 8 + 8 + 15) & ~(size_t)15;
            mpdu_cap += (size_t)b.desc[(size_t)p].trellis / 32 + 1; /* one A-MPDU delimiter per 32 bits at most */
        }
        b.psdu_off.assign((size_t)npkt, 0);
        if (slot < 0) {
            std::fprintf(stderr, "ieee80211 decode: no free GPU slot, batch dropped\n");
        } else if (b.bytes.reserve(bytes_cap, 0) && b.mpdus.reserve(mpdu_cap * sizeof(w80211_mpdu), 0)) {
            std::memset(&b.out, 0, sizeof(b.out));
            b.out.mem = W80211_MEM_HOST;
            b.out.bytes = b.bytes.data();
            b.out.bytes_capacity = b.bytes.capacity();
            b.out.psdu_off = b.psdu_off.data();
            b.out.mpdus = reinterpret_cast<w80211_mpdu*>(b.mpdus.data());
            b.out.mpdu_capacity = (uint32_t)std::min<size_t>(b.mpdus.capacity() / sizeof(w80211_mpdu), 0xffffffffu);
            std::memset(&b.in, 0, sizeof(b.in));
            b.in.npkt = npkt;
            b.in.desc = b.desc.data();
            if (b.lease) {
                b.in.llr_mem = W80211_MEM_LEASE;
                b.in.lease = b.lease;
            } else {
                b.in.llr_mem = W80211_MEM_HOST;
                b.in.llr = b.llr.data();
                b.in.nllr = partial ? (uint64_t)b.desc[(size_t)npkt].llr_off : b.fill;
            }
            if (w80211_decode_submit(d_ctx, slot, &b.in, &b.out) == W80211_OK) {
                b.slot = slot;
                d_slot_busy[slot] = true;
                sent = true;
            } else {
                std::fprintf(stderr, "ieee80211 decode: w80211_decode_submit: %s; batch dropped\n", w80211_last_error(d_ctx));
            }
        } else {
            std::fprintf(stderr, "ieee80211 decode: out of host memory for the results, batch dropped\n");
        }
    }
    b.state = sent ? SET_INFLIGHT : SET_READY;
    d_queue.push_back(idx);
    d_cur = -1;
    if (partial) d_partial_in = idx;
    take_free_set();
}

/* The oldest submitted batch: collected (waiting for it if asked to), its frames published, its set freed. */
bool decode_impl::finish_front(bool wait)
{
    if (d_queue.empty()) return false;
    const int idx = d_queue.front();
    batch_set& b = d_sets[idx];
    if (b.state == SET_INFLIGHT) {
        if (!wait && w80211_rx_poll(d_ctx, b.slot) == 0) return false;
        const double t_w = d_trace ? trace_now() : 0.0;
        const int r = w80211_rx_collect(d_ctx, b.slot);
        if (d_trace) d_t_wait += trace_now() - t_w;
        d_slot_busy[b.slot] = false;
        b.ok = r == W80211_OK;
        if (!b.ok) std::fprintf(stderr, "ieee80211 decode: w80211_decode_submit / w80211_rx_collect: %s; batch dropped\n", w80211_last_error(d_ctx));
        b.state = SET_READY;
    }
    if (b.ok) {
        const double t_p = d_trace ? trace_now() : 0.0;
        const w80211_mpdu* m = b.out.mpdus;
        for (uint32_t i = 0; i < b.out.n_mpdu; i++) publish(b.desc[m[i].pkt], b.bytes.data() + m[i].off, m[i].len, m[i].crc_ok != 0);
        if (d_trace) d_t_publish += trace_now() - t_p;
    }
    d_queue.pop_front();
    if (d_partial_in == idx) {
        /* it still holds the LLRs of the packet being received (no other set was free): it carries on as the set being filled */
        w80211_desc last = b.desc.back();
        const uint64_t have = b.fill - (uint64_t)last.llr_off;
        if (have) std::memmove(b.llr.data(), b.llr.data() + last.llr_off, (size_t)have * sizeof(float));
        last.llr_off = 0;
        b.desc.assign(1, last);
        b.fill = have;
        b.ncomplete = 0;
        b.lease = 0;
        b.state = SET_FILLING;
        d_partial_in = -1;
        d_cur = idx;
    } else {
        b.state = SET_FREE;
    }
    return true;
}

void decode_impl::drain_all()
{
    while (!d_queue.empty()) finish_front(true);
}

int decode_impl::general_work(int /*noutput_items*/,
                              gr_vector_int& ninput_items,
                              gr_vector_const_void_star& input_items,
                              gr_vector_void_star& /*output_items*/)
{
    const float* in = static_cast<const float*>(input_items[0]);
    const int navail = ninput_items[0];
    int consumed = 0;
    bool published = false;
    d_n_work++;
    const double t_i = d_trace ? trace_now() : 0.0;
    struct at_exit {
        decode_impl* c;
        double t;
        ~at_exit()
        {
            if (c->d_trace) c->d_t_ingest += trace_now() - t; /* the whole call; submit / wait / publish are counted again inside */
        }
    } guard{ this, t_i };
    while (finish_front(false)) published = true;
    take_free_set();
    while (d_cur >= 0) {
        bool stuck = false;
        while (consumed < navail && d_cur >= 0 && d_sets[d_cur].ncomplete < d_env.max_batch) {
            if (d_left == 0) {
                /* packet boundary: the reference idles until the first item carries tags (decode_impl.cc:68-70) */
                if (start_packet(nitems_read(0) + (uint64_t)consumed) != START_OK) {
                    stuck = true;
                    break;
                }
                if (d_cur < 0) break;
            }
            batch_set& b = d_sets[d_cur];
            const long n = std::min<long>(navail - consumed, d_left);
            const long ncopy = std::min(n, d_copy_left);
            if (ncopy > 0) {
                stream_copy(b.llr.data() + b.fill, in + consumed, (size_t)ncopy * sizeof(float));
                b.fill += (uint64_t)ncopy;
                d_copy_left -= ncopy;
                if (d_copy_left == 0) b.ncomplete++;
            }
            d_left -= n;
            consumed += (int)n;
        }
        if (d_cur < 0) break;
        batch_set& b = d_sets[d_cur];
        if (b.ncomplete == 0) break;
        const bool drained = consumed == navail || stuck;
        const double waited = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - b.oldest).count();
        if (b.ncomplete >= d_env.max_batch || (drained && b.ncomplete >= d_env.min_batch) || waited >= d_env.max_wait_ms)
            submit_current();
        else
            break;
        if (drained) break;
    }
    if (consumed == 0 && !published && !d_queue.empty()) {
        /* nothing else to do: wait for the oldest batch */
        finish_front(true);
        take_free_set();
    }
    if (d_trace && consumed == 0 && navail > 0) {
        static int said = 0;
        if (said++ < 20) {
            std::vector<gr::tag_t> tg;
            get_tags_in_range(tg, 0, nitems_read(0), nitems_read(0) + 1);
            std::fprintf(stderr, "ieee80211 decode trace: nothing consumed of %d items at %llu: d_left %ld d_cur %d queue %zu partial_in %d states %d %d %d "
                                 "published %d tags-at-item %zu ncomplete %d\n",
                         navail, (unsigned long long)nitems_read(0), d_left, d_cur, d_queue.size(), d_partial_in, (int)d_sets[0].state,
                         (int)d_sets[1].state, (int)d_sets[2].state, (int)published, tg.size(), d_cur >= 0 ? d_sets[d_cur].ncomplete : -1);
        }
    }
    consume_each(consumed);
    return 0;
}

} // namespace ieee80211
} // namespace gr
