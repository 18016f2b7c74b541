/* detect_impl.h -- gr::ieee80211::detect on w80211_detect_stream() (include/w80211rx.h).
 *
 * general_work() copies the arriving samples behind the history it keeps (the samples the previous call left unexamined and
 * at least 79 before them: presiso's windows), and when W80211_DETECT_CHUNK samples are waiting -- or the input has run dry --
 * hands the buffer to the GPU: presiso, run edges, the trigger's plateau logic and the sync block's IDLE / SYNC walk with
 * their state carried from buffer to buffer (w80211_detect_state), one thread per trigger for the L-LTF autocorrelation and
 * the two CFO steps.  What comes back is a list of flags; the output stream is written from it: one byte per examined
 * sample, the reference's tags on the flagged items. */
#ifndef INCLUDED_IEEE80211_DETECT_IMPL_H
#define INCLUDED_IEEE80211_DETECT_IMPL_H

#include <deque>
#include <vector>

#include <gnuradio/ieee80211/detect.h>
#include "demod_core.h" /* staged_array, rx_env */

namespace gr {
namespace ieee80211 {

class detect_impl : public detect
{
public:
    detect_impl();
    ~detect_impl() override;

    void forecast(int noutput_items, gr_vector_int& ninput_items_required) override;
    int general_work(int noutput_items,
                     gr_vector_int& ninput_items,
                     gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override;

private:
    struct flag {
        uint64_t pos; /* item of the output stream */
        float rad, snr, rssi;
    };
    void process();

    rx_env d_env;
    w80211_ctx* d_ctx;
    bool d_failed;
    size_t d_chunk;            /* W80211_DETECT_CHUNK: samples examined per GPU call at most (default 1 << 20) */

    staged_array<float> d_x;   /* history + samples not yet examined, interleaved (re, im) */
    size_t d_have, d_lead;     /* samples in d_x; the first d_lead are history */
    bool d_fresh;              /* samples have arrived since the last GPU call */
    w80211_detect_state d_state;
    bool d_have_state;
    std::vector<w80211_sync> d_syncs;

    uint64_t d_examined;       /* samples examined so far = output items decided */
    uint64_t d_emitted;        /* output items written so far */
    std::deque<flag> d_flags;  /* flags at items >= d_emitted, ascending */
};

} // namespace ieee80211
} // namespace gr

#endif
