/* decode_impl.h -- gr::ieee80211::decode on the GPU batch library (replaces lib/decode_impl.h of the reference).
 *
 * The reference decodes one packet at a time inside general_work(): depuncture + 64-state Viterbi over
 * `trellis` steps, traceback, descramble, A-MPDU walk, CRC-32 (lib/decode_impl.cc:60-520).  Here the
 * LLRs of whole packets are staged in page-locked memory and decoded by w80211_decode_batch()
 * (include/w80211rx.h) many packets at a time; the messages leave in arrival order with the
 * reference's layout [format, len lo, len hi, MPDU bytes, mcs]. */
#ifndef INCLUDED_IEEE80211_DECODE_IMPL_H
#define INCLUDED_IEEE80211_DECODE_IMPL_H

#include <chrono>
#include <gnuradio/ieee80211/decode.h>
#include "demod_core.h" /* staged_array, rx_env */

namespace gr {
namespace ieee80211 {

class decode_impl : public decode
{
public:
    explicit decode_impl(bool ifdebug);
    ~decode_impl() override;

    void forecast(int noutput_items, gr_vector_int& ninput_items_required) override;
    int general_work(int noutput_items,
                     gr_vector_int& ninput_items,
                     gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override;

private:
    bool start_packet(uint64_t abs_offset);
    void run_batch();
    void publish(const w80211_desc& d, const uint8_t* mpdu, uint32_t len, bool crc_ok);
    void debug_line(const w80211_desc& d, bool crc_ok);

    const bool d_debug;
    rx_env d_env;
    w80211_ctx* d_ctx;
    bool d_failed;

    staged_array<float> d_llr;
    std::vector<w80211_desc> d_desc;
    uint64_t d_fill; /* floats staged */
    int d_ncomplete;
    long d_left, d_copy_left;
    std::chrono::steady_clock::time_point d_oldest;

    staged_array<unsigned char> d_bytes, d_mpdus;
    std::vector<uint32_t> d_psdu_off;
    std::vector<unsigned char> d_msg;

    /* the counters the reference prints when ifdebug is set (lib/decode_impl.h:87-90) */
    uint64_t d_nPktCorrect;
    uint64_t d_legacyMcsCount[8], d_htMcsCount[8], d_vhtMcsCount[10];
};

} // namespace ieee80211
} // namespace gr

#endif
