/* decode_impl.cc -- decoder block: stream and message protocol of the reference
 * (lib/decode_impl.cc:33-162), batch decode on the GPU through w80211_decode_batch(). */
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
      d_fill(0),
      d_ncomplete(0),
      d_left(0),
      d_copy_left(0),
      d_nPktCorrect(0)
{
    message_port_register_out(pmt::mp("out"));
    std::memset(d_legacyMcsCount, 0, sizeof(d_legacyMcsCount));
    std::memset(d_htMcsCount, 0, sizeof(d_htMcsCount));
    std::memset(d_vhtMcsCount, 0, sizeof(d_vhtMcsCount));
    set_tag_propagation_policy(block::TPP_DONT);
    w80211_cfg cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.device = d_env.device;
    cfg.max_packets = d_env.max_batch;
    cfg.nslots = 1;
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
    }
}

decode_impl::~decode_impl()
{
    if (d_ctx) w80211_destroy(d_ctx);
}

void decode_impl::forecast(int /*noutput_items*/, gr_vector_int& ninput_items_required)
{
    ninput_items_required[0] = d_ncomplete > 0 ? 0 : 1;
}

/* Tags on the first LLR of a packet, keys and defaults as at decode_impl.cc:70-89, gates as at :93-121. */
bool decode_impl::start_packet(uint64_t abs_offset)
{
    std::vector<gr::tag_t> tags;
    get_tags_in_range(tags, 0, abs_offset, abs_offset + 1);
    if (tags.empty()) return false;
    w80211_desc d;
    std::memset(&d, 0, sizeof(d));
    long format = 99999, len = 99999, total = 99999, cr = 99999, mcs = 99999, ampdu = 99999, trellis = 99999;
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
    }
    if (total <= 0) return false; /* malformed stream: nothing to skip over */
    d_left = total;
    d_copy_left = 0;
    if (len > W80211_B_MAX || trellis > W80211_T_MAX) return true; /* skipped by the reference, decode_impl.cc:93-97 */
    if (trellis == 0) {
        /* VHT NDP: forward the 2x1 channel of the tag as [20, len lo, len hi, 256 floats] (decode_impl.cc:100-121).
         * Earlier packets are decoded first so the publication order stays the arrival order. */
        if (d_ncomplete > 0) run_batch();
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
        return true;
    }
    if (trellis < 0 || cr < 0 || cr > 3 || len < 0) return true; /* cannot be decoded; skip its LLRs */
    d.status = W80211_ST_OK;
    d.format = (int32_t)format;
    d.mcs = (int32_t)mcs;
    d.len = (int32_t)len;
    d.cr = (int32_t)cr;
    d.ampdu = (int32_t)ampdu;
    d.trellis = (int32_t)trellis;
    d.total = (int32_t)total;
    d.llr_off = (int64_t)((d_fill + 3) & ~(uint64_t)3); /* 16-byte aligned: no repacking inside the library */
    if (!d_llr.reserve((size_t)d.llr_off + (size_t)total + 64, (size_t)d_fill)) {
        if (!d_failed) std::fprintf(stderr, "ieee80211 decode: out of host memory for staging, dropping packets\n");
        d_failed = true;
        return true;
    }
    d_fill = (uint64_t)d.llr_off;
    d_copy_left = total;
    if (d_ncomplete == 0) d_oldest = std::chrono::steady_clock::now();
    d_desc.push_back(d);
    return true;
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

void decode_impl::run_batch()
{
    const int npkt = d_ncomplete;
    const bool partial = (int)d_desc.size() > npkt; /* the packet still arriving stays staged */
    if (!d_failed && npkt > 0) {
        size_t bytes_cap = 64, mpdu_cap = 64;
        for (int p = 0; p < npkt; p++) {
            bytes_cap += (size_t)((d_desc[(size_t)p].trellis + 7) / 8 + 8 + 15) & ~(size_t)15;
            mpdu_cap += (size_t)d_desc[(size_t)p].trellis / 32 + 1; /* one A-MPDU delimiter per 32 bits at most */
        }
        d_psdu_off.assign((size_t)npkt, 0);
        if (d_bytes.reserve(bytes_cap, 0) && d_mpdus.reserve(mpdu_cap * sizeof(w80211_mpdu), 0)) {
            w80211_batch_out out;
            std::memset(&out, 0, sizeof(out));
            out.mem = W80211_MEM_HOST;
            out.bytes = d_bytes.data();
            out.bytes_capacity = d_bytes.capacity();
            out.psdu_off = d_psdu_off.data();
            out.mpdus = reinterpret_cast<w80211_mpdu*>(d_mpdus.data());
            out.mpdu_capacity = (uint32_t)std::min<size_t>(d_mpdus.capacity() / sizeof(w80211_mpdu), 0xffffffffu);
            uint64_t nllr = d_fill;
            if (partial) nllr = (uint64_t)d_desc[(size_t)npkt].llr_off;
            int r = w80211_decode_batch(d_ctx, npkt, d_desc.data(), d_llr.data(), nllr, W80211_MEM_HOST, &out);
            if (r != W80211_OK) {
                std::fprintf(stderr, "ieee80211 decode: w80211_decode_batch: %s; batch dropped\n", w80211_last_error(d_ctx));
            } else {
                const w80211_mpdu* m = out.mpdus;
                for (uint32_t i = 0; i < out.n_mpdu; i++)
                    publish(d_desc[m[i].pkt], d_bytes.data() + m[i].off, m[i].len, m[i].crc_ok != 0);
            }
        } else {
            std::fprintf(stderr, "ieee80211 decode: out of host memory for the results, batch dropped\n");
        }
    }
    if (partial) {
        w80211_desc last = d_desc.back();
        const uint64_t have = d_fill - (uint64_t)last.llr_off;
        if (have) std::memmove(d_llr.data(), d_llr.data() + last.llr_off, (size_t)have * sizeof(float));
        last.llr_off = 0;
        d_desc.assign(1, last);
        d_fill = have;
    } else {
        d_desc.clear();
        d_fill = 0;
    }
    d_ncomplete = 0;
}

int decode_impl::general_work(int /*noutput_items*/,
                              gr_vector_int& ninput_items,
                              gr_vector_const_void_star& input_items,
                              gr_vector_void_star& /*output_items*/)
{
    const float* in = static_cast<const float*>(input_items[0]);
    const int navail = ninput_items[0];
    int consumed = 0;
    while (consumed < navail && d_ncomplete < d_env.max_batch) {
        if (d_left == 0) {
            /* packet boundary: the reference idles until the first item carries tags (decode_impl.cc:68-70) */
            if (!start_packet(nitems_read(0) + (uint64_t)consumed)) break;
        }
        const long n = std::min<long>(navail - consumed, d_left);
        const long ncopy = std::min(n, d_copy_left);
        if (ncopy > 0) {
            std::memcpy(d_llr.data() + d_fill, in + consumed, (size_t)ncopy * sizeof(float));
            d_fill += (uint64_t)ncopy;
            d_copy_left -= ncopy;
            if (d_copy_left == 0) d_ncomplete++;
        }
        d_left -= n;
        consumed += (int)n;
    }
    if (d_ncomplete > 0) {
        const bool drained = consumed == navail;
        const double waited = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - d_oldest).count();
        if (d_ncomplete >= d_env.max_batch || (drained && d_ncomplete >= d_env.min_batch) || waited >= d_env.max_wait_ms)
            run_batch();
    }
    consume_each(consumed);
    return 0;
}

} // namespace ieee80211
} // namespace gr
