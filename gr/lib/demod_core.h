/* demod_core.h -- what gr::ieee80211::demod and demod2 share in this implementation.
 *
 * The reference blocks (lib/demod_impl.cc:59-342, lib/demod2_impl.cc:58-349) walk one packet at a
 * time through a state machine and demodulate symbol by symbol inside general_work().  Here
 * general_work() only moves data:
 *
 *   ingest   the packets arriving on the input(s) -- tag on the first item, nsamp samples, 320 pad
 *            items (lib/signal_impl.cc:135-152,194-201) -- are copied into page-locked staging arrays
 *            and described by one w80211_pkt_in each;
 *   batch    when enough whole packets are staged (or the input has run dry) they go to the GPU in
 *            one w80211_demod_batch() call (include/w80211rx.h), which returns the per-packet
 *            descriptors (the reference's c8p_mod + tag values) and the LLR stream;
 *   emit     the LLRs of the packets that pass the reference's gates are written to the output in
 *            arrival order, `total` floats each, with the reference's tags on the first item.
 *
 * The stream protocol on both sides is the reference's; only the chunking differs, which a GNU Radio
 * flowgraph cannot observe.
 */
#ifndef INCLUDED_IEEE80211_DEMOD_CORE_H
#define INCLUDED_IEEE80211_DEMOD_CORE_H

#include <chrono>
#include <complex>
#include <cstdint>
#include <string>
#include <vector>

#include <gnuradio/block.h>
#include "w80211rx.h"

namespace gr {
namespace ieee80211 {

/* growable page-locked array (falls back to pageable memory when pinning fails) */
template <class T>
class staged_array
{
public:
    staged_array() : d_p(nullptr), d_cap(0), d_pinned(false) {}
    ~staged_array() { release(); }
    staged_array(const staged_array&) = delete;
    staged_array& operator=(const staged_array&) = delete;
    T* data() { return d_p; }
    size_t capacity() const { return d_cap; }
    /* keeps the first `keep` elements */
    bool reserve(size_t n, size_t keep);
    void release();

private:
    T* d_p;
    size_t d_cap;
    bool d_pinned;
};

/* environment knobs shared by the three blocks */
struct rx_env {
    int device;        /* W80211_DEVICE: CUDA device ordinal (default 0) */
    int max_batch;     /* W80211_MAX_BATCH: packets per GPU call at most (default 2048) */
    int min_batch;     /* W80211_MIN_BATCH: packets to wait for while input keeps arriving (default 1) */
    double max_wait_ms; /* W80211_MAX_WAIT_MS: oldest staged packet is never held longer than this (default 5) */
    static rx_env read();
};

class demod_core
{
public:
    demod_core(gr::block* owner, int nant, const char* label, int mupos = 0);
    ~demod_core();

    void forecast(int noutput_items, gr_vector_int& ninput_items_required);
    int work(int noutput_items,
             gr_vector_int& ninput_items,
             gr_vector_const_void_star& input_items,
             gr_vector_void_star& output_items);
    bool flush_pending() const { return d_ncomplete > 0 || !d_emit.empty(); }

private:
    struct emit_item {
        int pkt;        /* index into d_desc */
        size_t written; /* floats of this packet already written to the output */
        int ndp;        /* >= 0: a sounding packet, index into d_ndp_chan; its 1024 output items carry no LLRs */
    };

    int ingest(int navail, gr_vector_const_void_star& input_items);
    bool start_packet(uint64_t abs_offset);
    void run_batch();
    int emit(float* out, int room, int already);
    void write_tags(const w80211_desc& d, uint64_t abs_offset, int ndp);
    void report(const char* what);

    gr::block* d_blk;
    const int d_nant;
    const int d_mupos;
    const std::string d_label;
    rx_env d_env;
    w80211_ctx* d_ctx;
    bool d_failed; /* no usable GPU / library error: logged once, packets are dropped from then on */

    /* staging of the packets not yet sent to the GPU */
    staged_array<float> d_iq[2];       /* interleaved (re,im) */
    staged_array<float> d_chan;        /* [npkt][64][2] */
    std::vector<w80211_pkt_in> d_pkts;
    uint64_t d_fill;                   /* complex samples staged per antenna */
    int d_ncomplete;                   /* packets whose nsamp samples are all staged */
    long d_left, d_copy_left;          /* of the packet being received: items to consume / samples to copy */
    std::chrono::steady_clock::time_point d_oldest;

    /* results of the last batch, being written to the output */
    staged_array<float> d_llr;
    std::vector<w80211_desc> d_desc;
    std::vector<emit_item> d_emit;
    size_t d_emit_pos;
    std::vector<std::vector<std::complex<float>>> d_ndp_chan; /* tag "mu2x1chan" of the batch's sounding packets */
};

} // namespace ieee80211
} // namespace gr

#endif
