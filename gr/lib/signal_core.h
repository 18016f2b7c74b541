/* signal_core.h -- what gr::ieee80211::signal and signal2 share in this implementation.
 *
 * The reference blocks (lib/signal_impl.cc:62-212, lib/signal2_impl.cc:62-222) are a four-state machine
 * (TRIGGER / DEMOD / COPY / PAD) that handles one sync flag at a time with scalar cosf/sinf, three FFTs and
 * a 24-step Viterbi on the scheduler thread.  Here general_work() only moves data:
 *
 *   ingest   raw samples of the window are appended to page-locked staging arrays and every non-zero
 *            sync item becomes one w80211_sync (position, tags rad/snr/rssi);
 *   batch    w80211_signal_batch() (include/w80211rx.h) settles every flag whose packet lies wholly in the
 *            staged samples -- L-SIG demodulation and check for all flags in parallel, the state machine's
 *            skip rules on the host, CFO-rotated copy of the accepted packets -- and says where the next
 *            call has to resume; samples before that point are dropped from the staging;
 *   emit     accepted packets go to the output(s) in order: the reference's tags on the first item of
 *            output 0, nsamp samples, 320 pad items.
 *
 * The stream protocol on both sides is the reference's; only the chunking differs.
 */
#ifndef INCLUDED_IEEE80211_SIGNAL_CORE_H
#define INCLUDED_IEEE80211_SIGNAL_CORE_H

#include <cstdint>
#include <string>
#include <vector>

#include <gnuradio/block.h>
#include "demod_core.h"
#include "w80211rx.h"

namespace gr {
namespace ieee80211 {

class signal_core
{
public:
    signal_core(gr::block* owner, int nant, const char* label);
    ~signal_core();

    void forecast(int noutput_items, gr_vector_int& ninput_items_required);
    int work(int noutput_items,
             gr_vector_int& ninput_items,
             gr_vector_const_void_star& input_items,
             gr_vector_void_star& output_items);

private:
    void ingest(int n, gr_vector_const_void_star& input_items);
    void run_batch();
    int emit(gr_vector_void_star& output_items, int room, int already);
    void write_tags(const w80211_pkt_in& pk, const float* chan, long seq, uint64_t abs_offset);

    gr::block* d_blk;
    const int d_nant;
    const std::string d_label;
    rx_env d_env;
    w80211_ctx* d_ctx;
    bool d_failed;

    /* staged input: samples [d_base, d_base + d_fill) of the stream, flags not yet settled */
    staged_array<float> d_raw[2];
    uint64_t d_fill;
    uint64_t d_start;              /* trigger search position inside the staging */
    uint64_t d_need;               /* staged length at which the next batch call can make progress */
    std::vector<w80211_sync> d_syncs; /* pos relative to the staging */

    /* results of the last batch, being written to the output */
    staged_array<float> d_out[2];
    staged_array<float> d_chan;
    std::vector<w80211_pkt_in> d_pkts;
    size_t d_emit_pos;             /* packet being written */
    long d_emit_written;           /* of it: items already written (samples, then pad) */
    int d_seq;                     /* tag "seq": counter of accepted packets, wraps at 1e9 (signal_impl.cc:132-133) */
};

} // namespace ieee80211
} // namespace gr

#endif
