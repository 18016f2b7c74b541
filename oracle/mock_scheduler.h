/* Single-threaded stand-in for the GNU Radio scheduler, enough to drive one
 * gr::block (reference or replacement) over finite streams.
 *
 * It mimics what the real scheduler guarantees the blocks on this path:
 *   - general_work() is called from one thread, never re-entered;
 *   - each call sees a window of the unread input (<= chunk_in items) and room
 *     for chunk_out output items; if the block makes no progress the window
 *     grows, like GR honouring forecast(), until all remaining input is visible;
 *   - tags are visible through get_tags_in_range() with absolute offsets;
 *   - output items returned are appended to the output stream;
 *   - with honor_forecast, noutput_items is halved until forecast() is satisfied by the visible input.
 * The run ends when the block makes no progress for MOCK_IDLE_STOP calls in a
 * row while seeing all the remaining input (the reference state machines take
 * up to four zero-progress calls between states).
 * Test infrastructure only. */
#pragma once
#include <algorithm>
#include <gnuradio/block.h>

namespace gr {

enum { MOCK_IDLE_GROW = 6, MOCK_IDLE_STOP = 12 };

inline long mock_run(block& b,
                     const std::vector<const mock_stream*>& ins,
                     const std::vector<mock_stream*>& outs,
                     long chunk_in,
                     long chunk_out,
                     bool honor_forecast = false)
{
    const size_t nin = ins.size(), nout = outs.size();
    b.mock_in = ins;
    b.mock_out = outs;
    b.mock_nread.assign(nin, 0);
    b.mock_nwritten.assign(nout, 0);
    for (size_t o = 0; o < nout; o++) b.mock_nwritten[o] = outs[o]->nitems();
    b.mock_consumed.assign(nin, 0);

    std::vector<uint64_t> pos(nin, 0);
    if (chunk_in <= 0) chunk_in = 1L << 40;
    if (chunk_out <= 0) chunk_out = 8192;
    long win_in = chunk_in, win_out = chunk_out;
    int idle = 0;
    long calls = 0;
    std::vector<std::vector<uint8_t>> scratch(nout);

    for (;;) {
        gr_vector_int ninput(nin);
        gr_vector_const_void_star in_ptrs(nin);
        gr_vector_void_star out_ptrs(nout);
        bool sees_all = true;
        for (size_t i = 0; i < nin; i++) {
            uint64_t rem = ins[i]->nitems() - pos[i];
            uint64_t n = std::min<uint64_t>(rem, (uint64_t)win_in);
            if (n < rem) sees_all = false;
            ninput[i] = (int)std::min<uint64_t>(n, 0x7fffffffu);
            in_ptrs[i] = ins[i]->data.data() + pos[i] * ins[i]->itemsize;
        }
        for (size_t o = 0; o < nout; o++) {
            scratch[o].assign((size_t)win_out * outs[o]->itemsize, 0);
            out_ptrs[o] = scratch[o].data();
        }
        std::fill(b.mock_consumed.begin(), b.mock_consumed.end(), 0);
        long call_out = win_out;
        if (honor_forecast) {
            /* like GR's single-threaded executor: shrink noutput_items until forecast() fits the visible input; blocks
             * that index their input by noutput_items (trigger, sync) depend on it */
            for (; call_out > 0; call_out /= 2) {
                gr_vector_int need(nin, 0);
                b.forecast((int)call_out, need);
                bool fits = true;
                for (size_t i = 0; i < nin; i++) fits = fits && need[i] <= ninput[i];
                if (fits) break;
            }
            if (call_out == 0) break; /* starved: the flowgraph would end here */
        }
        int produced = b.general_work((int)call_out, ninput, in_ptrs, out_ptrs);
        calls++;
        bool progress = false;
        if (produced > 0) {
            progress = true;
            for (size_t o = 0; o < nout; o++) {
                outs[o]->data.insert(outs[o]->data.end(),
                                     scratch[o].begin(),
                                     scratch[o].begin() + (size_t)produced * outs[o]->itemsize);
                b.mock_nwritten[o] += produced;
            }
        }
        for (size_t i = 0; i < nin; i++) {
            if (b.mock_consumed[i] > 0) {
                progress = true;
                pos[i] += b.mock_consumed[i];
                b.mock_nread[i] += b.mock_consumed[i];
            }
        }
        if (progress) {
            idle = 0;
            win_in = chunk_in;
            win_out = chunk_out;
            continue;
        }
        idle++;
        if (idle >= MOCK_IDLE_GROW && (!sees_all || win_out < 4096)) {
            if (!sees_all) win_in *= 2;
            if (win_out < 4096) win_out *= 2; /* largest atomic output: 1024 (NDP) / 832 LLRs */
            idle = 0;
            continue;
        }
        if (idle >= MOCK_IDLE_STOP) break;
    }
    return calls;
}

} // namespace gr
