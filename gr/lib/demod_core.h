/* demod_core.h -- what gr::ieee80211::demod and demod2 share in this implementation.
 *
 * The reference blocks (lib/demod_impl.cc:59-342, lib/demod2_impl.cc:58-349) walk one packet at a
 * time through a state machine and demodulate symbol by symbol inside general_work().  Here
 * general_work() only moves data, and never waits for the GPU while it has anything else to do:
 *
 *   ingest   the packets arriving on the input(s) -- tag on the first item, nsamp samples, 320 pad
 *            items (lib/signal_impl.cc:135-152,194-201) -- are copied into the page-locked staging arrays
 *            of the batch being filled and described by one w80211_pkt_in each;
 *   submit   when enough whole packets are staged (or the input has run dry) the batch goes to the GPU with
 *            w80211_demod_submit() (include/w80211rx.h), which returns at once; ingest continues into the
 *            next of three batch sets while up to two batches are in flight;
 *   emit     finished batches are collected in submission order and the LLRs of the packets that pass the
 *            reference's gates are written to the output in arrival order, `total` floats each, with the
 *            reference's tags on the first item.
 *
 * LLR hand-off.  The reference's demod -> decode edge carries 43 floats per payload byte.  When this block
 * feeds this repo's decode block in the same process the LLRs can stay in HBM: the batch's LLR buffer is
 * exported as a lease (w80211_llr_export) and every packet's first item carries two more tags,
 * "w80211_llr" (the lease handle) and "w80211_off" (the packet's offset in it).  W80211_LLR_HANDOFF picks:
 *   host    LLR items only, exactly the reference's stream;
 *   both    LLR items and the lease tags (decode skips its staging copy and its upload);
 *   device  lease tags only: the `total` items of a packet are produced but NOT written (like the 1024 items of
 *           an NDP in the reference); only a decode block of this repo can use such a stream;
 *   auto    (default) `device` while a decode block of this repo lives in the process, else `host`.  A flowgraph
 *           that also taps demod's output (file sink, scope) sets `both` or `host`.
 *
 * The stream protocol on both sides is the reference's; only the chunking differs, which a GNU Radio
 * flowgraph cannot observe.
 */
#ifndef INCLUDED_IEEE80211_DEMOD_CORE_H
#define INCLUDED_IEEE80211_DEMOD_CORE_H

#include <chrono>
#include <complex>
#include <cstdint>
#include <deque>
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

/* copy into page-locked staging that the CPU will not read again (non-temporal stores where the CPU has AVX2) */
void stream_copy(void* dst, const void* src, size_t bytes);

enum llr_handoff { HANDOFF_AUTO = 0, HANDOFF_HOST, HANDOFF_BOTH, HANDOFF_DEVICE };

/* environment knobs shared by the three blocks */
struct rx_env {
    int device;        /* W80211_DEVICE: CUDA device ordinal (default 0) */
    int max_batch;     /* W80211_MAX_BATCH: packets per GPU call at most (default 2048) */
    int min_batch;     /* W80211_MIN_BATCH: packets to wait for while input keeps arriving (default 1) */
    double max_wait_ms; /* W80211_MAX_WAIT_MS: oldest staged packet is never held longer than this (default 5) */
    int handoff;       /* W80211_LLR_HANDOFF: auto | host | both | device (see above) */
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
    /* true while staged packets or unfinished batches can still turn into output without new input */
    bool flush_pending() const;

private:
    enum { NSETS = 3, NSLOTS = 2 };
    struct emit_item {
        int pkt;        /* index into desc */
        size_t written; /* floats of this packet already produced */
        int ndp;        /* >= 0: a sounding packet, index into ndp_chan; its 1024 output items carry no LLRs */
    };
    enum set_state { SET_FREE, SET_FILLING, SET_INFLIGHT, SET_READY };
    /* one batch: staging while it fills, the call's argument blocks while it is in flight, results while it is emitted */
    struct batch_set {
        set_state state = SET_FREE;
        staged_array<float> iq[2]; /* interleaved (re,im) */
        staged_array<float> chan;  /* [npkt][64][2] */
        staged_array<float> llr;   /* results, when the LLR items are wanted on the host */
        std::vector<w80211_pkt_in> pkts;
        uint64_t fill = 0;  /* complex samples staged per antenna */
        int ncomplete = 0;  /* packets whose nsamp samples are all staged */
        std::chrono::steady_clock::time_point oldest;
        int slot = -1;
        int npkt = 0;       /* packets submitted (the one still arriving is not among them) */
        bool items_on_host = true;
        bool want_lease = false;
        uint64_t lease = 0;
        w80211_batch_in in;
        w80211_batch_out out;
        std::vector<w80211_desc> desc;
        std::vector<emit_item> emit;
        size_t emit_pos = 0;
        std::vector<std::vector<std::complex<float>>> ndp_chan; /* tag "mu2x1chan" of the batch's sounding packets */
    };

    int ingest(int base, int navail, gr_vector_const_void_star& input_items);
    bool start_packet(uint64_t abs_offset);
    bool take_free_set();
    void submit_current();
    bool collect_set(int idx, bool wait);
    int emit(batch_set& b, float* out, int room, int already);
    void write_tags(const batch_set& b, const w80211_desc& d, uint64_t abs_offset, int ndp);
    void report(const char* what);

    gr::block* d_blk;
    const int d_nant;
    const int d_mupos;
    const std::string d_label;
    rx_env d_env;
    w80211_ctx* d_ctx;
    bool d_failed; /* no usable GPU / library error: logged once, packets are dropped from then on */

    batch_set d_sets[NSETS];
    int d_cur;               /* set being filled, -1: none free at the moment */
    int d_partial_in;        /* submitted set that still holds the samples of the packet being received, -1: none */
    std::deque<int> d_queue; /* submitted sets, oldest first (INFLIGHT or READY) */
    bool d_slot_busy[NSLOTS];
    long d_left, d_copy_left; /* of the packet being received: items to consume / samples to copy */

    /* W80211_TRACE=1: where general_work() spends its time, printed when the block is destroyed */
    bool d_trace;
    double d_t_ingest, d_t_submit, d_t_wait, d_t_emit;
    uint64_t d_n_work, d_n_batches, d_n_pkts;
};

/* seconds since an arbitrary origin, for the W80211_TRACE accounting of the blocks */
inline double trace_now()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

} // namespace ieee80211
} // namespace gr

#endif
