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
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <exception>
#include <mutex>
#include <thread>
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

/* Two blocks joined by one bounded edge (demod -> decode), run in turn by one thread the way GNU Radio's single-threaded
 * scheduler would: the upstream block writes straight into the edge's buffer (`ring_items` items; nothing is zero-filled
 * or appended, so items a block produces without writing them cost nothing), the downstream block reads from it, and the
 * buffer is compacted when its tail is reached.  Tags of the edge are dropped once consumed.  Returns the number of
 * general_work calls; the downstream block's messages stay in its mock_msgs. */
inline long mock_run_chain(block& up,
                           block& down,
                           const std::vector<const mock_stream*>& ins,
                           long ring_items,
                           long chunk_in,
                           long chunk_out)
{
    const size_t nin = ins.size();
    mock_stream mid(up.mock_outsig ? up.mock_outsig->sizeof_stream_item(0) : 4);
    mid.tags_sorted = true;
    if (ring_items < 4 * 4096) ring_items = 4 * 4096;
    mid.data.resize((size_t)ring_items * mid.itemsize);
    up.mock_in = ins;
    up.mock_out.assign(1, &mid);
    up.mock_nread.assign(nin, 0);
    up.mock_nwritten.assign(1, 0);
    up.mock_consumed.assign(nin, 0);
    down.mock_in.assign(1, &mid);
    down.mock_out.clear();
    down.mock_nread.assign(1, 0);
    down.mock_nwritten.clear();
    down.mock_consumed.assign(1, 0);
    std::vector<uint64_t> pos(nin, 0);
    uint64_t w = 0, r = 0, base = 0; /* absolute item indices: written, read, and of mid.data[0] */
    if (chunk_in <= 0) chunk_in = 1L << 40;
    if (chunk_out <= 0) chunk_out = 8192;
    long win_in = chunk_in;
    int idle = 0;
    long calls = 0;
    for (;;) {
        bool progress = false;
        /* ---- upstream */
        if ((long)(w - base) + chunk_out > ring_items && r > base) {
            const uint64_t keep = w - r;
            if (keep) std::memmove(mid.data.data(), mid.data.data() + (size_t)(r - base) * mid.itemsize, (size_t)keep * mid.itemsize);
            base = r;
        }
        long room = std::min<long>(chunk_out, ring_items - (long)(w - base));
        bool sees_all = true;
        if (room > 0) {
            gr_vector_int ninput(nin);
            gr_vector_const_void_star in_ptrs(nin);
            gr_vector_void_star out_ptrs(1);
            for (size_t i = 0; i < nin; i++) {
                uint64_t rem = ins[i]->nitems() - pos[i];
                uint64_t n = std::min<uint64_t>(rem, (uint64_t)win_in);
                if (n < rem) sees_all = false;
                ninput[i] = (int)std::min<uint64_t>(n, 0x7fffffffu);
                in_ptrs[i] = ins[i]->data.data() + pos[i] * ins[i]->itemsize;
            }
            out_ptrs[0] = mid.data.data() + (size_t)(w - base) * mid.itemsize;
            std::fill(up.mock_consumed.begin(), up.mock_consumed.end(), 0);
            const int produced = up.general_work((int)room, ninput, in_ptrs, out_ptrs);
            calls++;
            if (produced > 0) {
                w += (uint64_t)produced;
                up.mock_nwritten[0] = w;
                progress = true;
            }
            for (size_t i = 0; i < nin; i++)
                if (up.mock_consumed[i] > 0) {
                    pos[i] += up.mock_consumed[i];
                    up.mock_nread[i] += up.mock_consumed[i];
                    progress = true;
                }
        }
        /* ---- downstream */
        {
            gr_vector_int ninput(1, (int)std::min<uint64_t>(w - r, 0x7fffffffu));
            gr_vector_const_void_star in_ptrs(1, mid.data.data() + (size_t)(r - base) * mid.itemsize);
            gr_vector_void_star out_ptrs;
            down.mock_consumed[0] = 0;
            down.general_work(0, ninput, in_ptrs, out_ptrs);
            calls++;
            if (down.mock_consumed[0] > 0) {
                r += (uint64_t)down.mock_consumed[0];
                down.mock_nread[0] = r;
                progress = true;
                size_t k = 0;
                while (k < mid.tags.size() && mid.tags[k].offset < r) k++;
                if (k > 4096 || k == mid.tags.size()) mid.tags.erase(mid.tags.begin(), mid.tags.begin() + (long)k);
            }
        }
        if (progress) {
            idle = 0;
            win_in = chunk_in;
            continue;
        }
        idle++;
        if (idle >= MOCK_IDLE_GROW && !sees_all) {
            win_in *= 2;
            idle = 0;
            continue;
        }
        if (idle >= MOCK_IDLE_STOP) break;
    }
    return calls;
}

/* The same two blocks, each on a thread of its own like under GNU Radio's thread-per-block scheduler.  The edge is the same
 * linear buffer of `ring_items` items: the upstream thread writes behind the last written item while the downstream thread
 * reads its window; when the tail is reached the unread items are moved to the front, which the upstream thread does while
 * holding the lock the downstream thread holds during its general_work() (windows stay contiguous, as windows of GNU Radio's
 * double-mapped buffers are).  Positions and the edge's tags are guarded by one mutex; a thread that made no progress waits
 * for the other on a condition variable (0.1 ms at most).  The upstream thread ends after it has seen all its input and made no
 * progress MOCK_IDLE_STOP times in a row, the downstream one after that and the same.  Returns the general_work calls of both. */
inline long mock_run_chain_threads(block& up, block& down, const std::vector<const mock_stream*>& ins, long ring_items, long chunk_in,
                                   long chunk_out)
{
    const size_t nin = ins.size();
    mock_stream mid(up.mock_outsig ? up.mock_outsig->sizeof_stream_item(0) : 4);
    mid.tags_sorted = true;
    std::mutex mu, window_mu;
    std::condition_variable cv;
    mid.lock = &mu;
    if (ring_items < 4 * 4096) ring_items = 4 * 4096;
    if (chunk_in <= 0) chunk_in = 1L << 40;
    if (chunk_out <= 0) chunk_out = 8192;
    if (chunk_out > ring_items / 2) chunk_out = ring_items / 2;
    mid.data.resize((size_t)ring_items * mid.itemsize);
    up.mock_in = ins;
    up.mock_out.assign(1, &mid);
    up.mock_nread.assign(nin, 0);
    up.mock_nwritten.assign(1, 0);
    up.mock_consumed.assign(nin, 0);
    down.mock_in.assign(1, &mid);
    down.mock_out.clear();
    down.mock_nread.assign(1, 0);
    down.mock_nwritten.clear();
    down.mock_consumed.assign(1, 0);
    uint64_t w = 0, r = 0, base = 0; /* absolute items written / read / at data[0]; guarded by mu */
    bool up_done = false;
    std::atomic<long> calls(0);
    /* a flowgraph that stops moving is a bug to report, not to wait out: no progress on either side for MOCK_STALL_S seconds
     * ends the run with a line on stderr */
    const double MOCK_STALL_S = 20.0;
    auto now_s = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    std::atomic<double> last_progress(now_s());
    std::atomic<bool> stalled(false);
    /* a thread notifies only when the other one sleeps: a wake-up per call would cost more than the blocks' work */
    std::atomic<int> sleepers(0);
    auto nap = [&](std::unique_lock<std::mutex>& lk) {
        sleepers++;
        cv.wait_for(lk, std::chrono::microseconds(100)); /* short: the end of a run is 2 x MOCK_IDLE_STOP of these */
        sleepers--;
    };
    auto wake = [&] {
        if (sleepers.load() > 0) cv.notify_all();
    };

    std::exception_ptr up_error;
    std::thread producer([&] {
      try {
        std::vector<uint64_t> pos(nin, 0);
        long win_in = chunk_in;
        int idle = 0;
        for (;;) {
            long room;
            uint64_t at;
            {
                std::unique_lock<std::mutex> lk(mu);
                if ((long)(w - base) + chunk_out > ring_items && r > base) {
                    lk.unlock();
                    std::lock_guard<std::mutex> wl(window_mu); /* the reader is between two calls */
                    lk.lock();
                    const uint64_t keep = w - r;
                    if (keep) std::memmove(mid.data.data(), mid.data.data() + (size_t)(r - base) * mid.itemsize, (size_t)keep * mid.itemsize);
                    base = r;
                }
                room = std::min<long>(chunk_out, ring_items - (long)(w - base));
                at = w - base;
            }
            bool progress = false, sees_all = true;
            if (room > 0) {
                gr_vector_int ninput(nin);
                gr_vector_const_void_star in_ptrs(nin);
                gr_vector_void_star out_ptrs(1);
                for (size_t i = 0; i < nin; i++) {
                    const uint64_t rem = ins[i]->nitems() - pos[i];
                    const uint64_t n = std::min<uint64_t>(rem, (uint64_t)win_in);
                    if (n < rem) sees_all = false;
                    ninput[i] = (int)std::min<uint64_t>(n, 0x7fffffffu);
                    in_ptrs[i] = ins[i]->data.data() + pos[i] * ins[i]->itemsize;
                }
                out_ptrs[0] = mid.data.data() + (size_t)at * mid.itemsize; /* only this thread moves w and base */
                std::fill(up.mock_consumed.begin(), up.mock_consumed.end(), 0);
                const int produced = up.general_work((int)room, ninput, in_ptrs, out_ptrs);
                calls++;
                for (size_t i = 0; i < nin; i++)
                    if (up.mock_consumed[i] > 0) {
                        pos[i] += up.mock_consumed[i];
                        up.mock_nread[i] += up.mock_consumed[i];
                        progress = true;
                    }
                if (produced > 0) {
                    std::unique_lock<std::mutex> lk(mu);
                    w += (uint64_t)produced;
                    up.mock_nwritten[0] = w;
                    progress = true;
                    wake();
                }
            }
            if (progress) {
                idle = 0;
                win_in = chunk_in;
                last_progress = now_s();
                continue;
            }
            idle++;
            if (idle >= MOCK_IDLE_GROW && !sees_all && room > 0) {
                win_in *= 2;
                idle = 0;
                continue;
            }
            if (idle >= MOCK_IDLE_STOP && sees_all && room > 0) break;
            if (stalled || now_s() - last_progress > MOCK_STALL_S) {
                std::unique_lock<std::mutex> lk(mu);
                if (!stalled)
                    fprintf(stderr, "mock_run_chain_threads: STALLED (upstream): w %llu r %llu base %llu room %ld idle %d sees_all %d\n",
                            (unsigned long long)w, (unsigned long long)r, (unsigned long long)base, room, idle, (int)sees_all);
                stalled = true;
                break;
            }
            std::unique_lock<std::mutex> lk(mu);
            nap(lk);
        }
      } catch (...) {
        up_error = std::current_exception(); /* rethrown by the calling thread after the join */
      }
        std::unique_lock<std::mutex> lk(mu);
        up_done = true;
        cv.notify_all();
    });

    int idle = 0;
    std::exception_ptr down_error;
    try {
    for (;;) {
        bool done, moved = false;
        {
            std::lock_guard<std::mutex> wl(window_mu); /* the window must not move while the block reads it */
            uint64_t avail, at;
            {
                std::unique_lock<std::mutex> lk(mu);
                avail = w - r;
                at = r - base;
                done = up_done;
            }
            gr_vector_int ninput(1, (int)std::min<uint64_t>(avail, 0x7fffffffu));
            gr_vector_const_void_star in_ptrs(1, mid.data.data() + (size_t)at * mid.itemsize);
            gr_vector_void_star out_ptrs;
            down.mock_consumed[0] = 0;
            down.general_work(0, ninput, in_ptrs, out_ptrs);
            calls++;
            if (down.mock_consumed[0] > 0) {
                std::unique_lock<std::mutex> lk(mu);
                r += (uint64_t)down.mock_consumed[0];
                down.mock_nread[0] = r;
                size_t k = 0;
                while (k < mid.tags.size() && mid.tags[k].offset < r) k++;
                if (k > 4096 || k == mid.tags.size()) mid.tags.erase(mid.tags.begin(), mid.tags.begin() + (long)k);
                wake();
                moved = true;
            }
        }
        if (moved) {
            idle = 0;
            last_progress = now_s();
            continue;
        }
        if (done && ++idle >= MOCK_IDLE_STOP) break;
        if (stalled || now_s() - last_progress > MOCK_STALL_S) {
            std::unique_lock<std::mutex> lk(mu);
            if (!stalled)
                fprintf(stderr, "mock_run_chain_threads: STALLED (downstream): w %llu r %llu base %llu done %d idle %d\n", (unsigned long long)w,
                        (unsigned long long)r, (unsigned long long)base, (int)done, idle);
            stalled = true;
            cv.notify_all();
            break;
        }
        std::unique_lock<std::mutex> lk(mu);
        nap(lk);
    }
    } catch (...) {
        down_error = std::current_exception();
        stalled = true; /* stops the upstream thread */
        cv.notify_all();
    }
    producer.join();
    mid.lock = nullptr;
    if (down_error) std::rethrow_exception(down_error);
    if (up_error) std::rethrow_exception(up_error);
    return stalled ? -2 : calls.load();
}

} // namespace gr
