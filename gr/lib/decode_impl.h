/* decode_impl.h -- gr::ieee80211::decode on the GPU batch library (replaces lib/decode_impl.h of the reference).
 *
 * The reference decodes one packet at a time inside general_work(): depuncture + 64-state Viterbi over
 * `trellis` steps, traceback, descramble, A-MPDU walk, CRC-32 (lib/decode_impl.cc:60-520).  Here the
 * packets of a batch are decoded together by w80211_decode_submit() (include/w80211rx.h), which does not
 * wait: while up to two batches are on the GPU the block keeps staging the next one (three batch sets), and
 * finished batches are collected in submission order, so the messages leave in arrival order with the
 * reference's layout [format, len lo, len hi, MPDU bytes, mcs].
 *
 * Where the LLRs come from.  A packet whose first item carries this repo's demod's tags "w80211_llr" /
 * "w80211_off" (gr/lib/demod_core.h) is decoded from the LLR lease they name, straight out of HBM: its `total`
 * stream items are consumed without being read.  Any other packet -- a foreign demod, W80211_LLR_HANDOFF=host,
 * a lease that has been retired in the meantime -- is staged from the stream items in page-locked memory and
 * uploaded.  A batch holds packets of one source only. */
#ifndef INCLUDED_IEEE80211_DECODE_IMPL_H
#define INCLUDED_IEEE80211_DECODE_IMPL_H

#include <chrono>
#include <deque>
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
    enum { NSETS = 3, NSLOTS = 2 };
    enum set_state { SET_FREE, SET_FILLING, SET_INFLIGHT, SET_READY };
    struct batch_set {
        set_state state = SET_FREE;
        staged_array<float> llr; /* host-sourced packets: their LLRs at 16-byte aligned offsets */
        std::vector<w80211_desc> desc;
        uint64_t fill = 0;       /* floats staged */
        int ncomplete = 0;
        uint64_t lease = 0;      /* != 0: every packet of the set reads this lease */
        std::chrono::steady_clock::time_point oldest;
        int slot = -1;
        int npkt = 0;
        bool ok = false;         /* the batch was decoded: out holds results */
        w80211_decode_in in;
        w80211_batch_out out;
        staged_array<unsigned char> bytes, mpdus;
        std::vector<uint32_t> psdu_off;
    };
    enum { START_NONE = 0, START_OK = 1, START_LATER = -1 };

    int start_packet(uint64_t abs_offset);
    bool take_free_set();
    void submit_current();
    bool finish_front(bool wait);
    void drain_all();
    void publish(const w80211_desc& d, const uint8_t* mpdu, uint32_t len, bool crc_ok);
    void debug_line(const w80211_desc& d, bool crc_ok);

    const bool d_debug;
    rx_env d_env;
    w80211_ctx* d_ctx;
    bool d_failed;

    batch_set d_sets[NSETS];
    int d_cur;               /* set being filled, -1: none free at the moment */
    int d_partial_in;        /* submitted set that still holds the LLRs of the packet being received */
    std::deque<int> d_queue; /* submitted sets, oldest first */
    bool d_slot_busy[NSLOTS];
    long d_left, d_copy_left;
    uint64_t d_lost_leases;  /* packets dropped because their LLRs existed only in a lease that had been retired */

    std::vector<unsigned char> d_msg;

    /* W80211_TRACE=1: where general_work() spends its time, printed when the block is destroyed */
    bool d_trace;
    double d_t_ingest, d_t_submit, d_t_wait, d_t_publish;
    uint64_t d_n_work, d_n_batches, d_n_pkts, d_n_leased;

    /* the counters the reference prints when ifdebug is set (lib/decode_impl.h:87-90) */
    uint64_t d_nPktCorrect;
    uint64_t d_legacyMcsCount[8], d_htMcsCount[8], d_vhtMcsCount[10];
};

} // namespace ieee80211
} // namespace gr

#endif
