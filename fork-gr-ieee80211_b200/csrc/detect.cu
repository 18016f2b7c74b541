/* detect.cu -- packet detection on the GPU: the data-parallel parts of the reference's trigger
 * block (lib/trigger_impl.cc:60-117) and sync block (lib/sync_impl.cc:61-196).
 *
 *   run_edges_kernel   one thread per sample: where runs of preac > 0.3 begin and end (the only thing the trigger's
 *                      plateau counter, its running maximum and its flags depend on).  The few edges go to two lists in
 *                      ascending position (count per block, scan, write at offset + rank: no sort, no library kernel).
 *   (host)             w80211_detect_batch walks the runs in order: plateau > 20 -> 80-sample countdown -> trigger flag,
 *                      one countdown at a time, one per run; then the sync block's IDLE / SYNC alternation (a trigger
 *                      inside the 111 samples a SYNC pass consumes is never seen).
 *   sync_event_kernel  one thread per surviving trigger: (a) the preconj sample at the last 0x02 flag the sync block saw
 *                      before the trigger = the first occurrence of the run maximum, (b) ltf_autoCorrelation with the
 *                      reference's sequential binary32 running sums over 111 offsets, first maximum, the 0.8 * max
 *                      plateau edges, their midpoint, (c) ltf_cfo: coarse step from preconj, 128 rotated samples, fine
 *                      step from the 64-lag product sum, (d) snr and rssi.
 * Every per-trigger result depends on at most a few hundred samples, so a thread per trigger keeps the reference's
 * operation order (bit-exact sums) and the batch supplies the parallelism.
 */
#include "common.cuh"

__device__ __forceinline__ bool above_thr(float v) { return v > 0.3f; }

/* Edges of the runs of preac > 0.3, listed in ascending position without a sort: a block covers EDGE_TILES x EDGE_BLOCK
 * consecutive samples; pass 1 (WRITE = false) counts the block's starts and ends, a one-block scan turns the counts into
 * offsets, pass 2 (WRITE = true) finds the same edges again and stores each at offset + its rank inside the block (ballot
 * prefix inside a warp, warp totals through shared memory) -- ranks follow positions, so both lists come out ordered. */
#define EDGE_BLOCK 1024
#define EDGE_TILES 4
template <bool WRITE>
__global__ void __launch_bounds__(EDGE_BLOCK) run_edges_kernel(const float* __restrict__ ac, unsigned long long n, int prev_above0,
                                                               uint2* __restrict__ blk_cnt, const uint2* __restrict__ blk_off,
                                                               unsigned int cap, unsigned int* __restrict__ starts,
                                                               unsigned int* __restrict__ ends, unsigned int* __restrict__ counters)
{
    __shared__ unsigned int wtot[2][32];
    __shared__ unsigned int run[2], tile_tot[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long base = (unsigned long long)blockIdx.x * (EDGE_BLOCK * EDGE_TILES);
    if (tid == 0) {
        run[0] = WRITE ? blk_off[blockIdx.x].x : 0u;
        run[1] = WRITE ? blk_off[blockIdx.x].y : 0u;
    }
    __syncthreads();
    for (int t = 0; t < EDGE_TILES; t++) {
        const unsigned long long i = base + (unsigned long long)t * EDGE_BLOCK + tid;
        bool is_s = false, is_e = false;
        if (i < n) {
            const bool a = above_thr(__ldg(ac + i));
            const bool b = i ? above_thr(__ldg(ac + i - 1)) : (prev_above0 != 0);
            is_s = a && !b;
            is_e = !a && b;
        }
        const unsigned int bs = __ballot_sync(0xffffffffu, is_s), be = __ballot_sync(0xffffffffu, is_e);
        if (lane == 0) {
            wtot[0][warp] = __popc(bs);
            wtot[1][warp] = __popc(be);
        }
        __syncthreads();
        if (warp == 0) { /* exclusive scan of the 32 warp totals, both lists; lane 31 ends up with the tile's totals */
            const unsigned int vs = wtot[0][lane], ve = wtot[1][lane];
            unsigned int xs = vs, xe = ve;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned int us = __shfl_up_sync(0xffffffffu, xs, d), ue = __shfl_up_sync(0xffffffffu, xe, d);
                if (lane >= d) {
                    xs += us;
                    xe += ue;
                }
            }
            wtot[0][lane] = xs - vs;
            wtot[1][lane] = xe - ve;
            if (lane == 31) {
                tile_tot[0] = xs;
                tile_tot[1] = xe;
            }
        }
        __syncthreads();
        if (WRITE) {
            if (is_s) {
                const unsigned int k = run[0] + wtot[0][warp] + __popc(bs & ((1u << lane) - 1u));
                if (k < cap) starts[k] = (unsigned int)i;
            }
            if (is_e) {
                const unsigned int k = run[1] + wtot[1][warp] + __popc(be & ((1u << lane) - 1u));
                if (k < cap) ends[k] = (unsigned int)i;
            }
        }
        __syncthreads();
        if (tid == 0) {
            run[0] += tile_tot[0];
            run[1] += tile_tot[1];
        }
    }
    __syncthreads();
    if (!WRITE && tid == 0) {
        blk_cnt[blockIdx.x] = make_uint2(run[0], run[1]);
        if (run[0]) atomicAdd(&counters[0], run[0]);
        if (run[1]) atomicAdd(&counters[1], run[1]);
    }
}

/* exclusive scan of the per-block (starts, ends) counts: one block, each thread a contiguous chunk */
__global__ void __launch_bounds__(1024) edges_scan_kernel(const uint2* __restrict__ cnt, uint2* __restrict__ off, int nblk)
{
    __shared__ unsigned int ws[2][32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (nblk + 1023) / 1024, lo = tid * per, hi = min(nblk, lo + per);
    unsigned int s0 = 0, s1 = 0;
    for (int k = lo; k < hi; k++) {
        const uint2 c = cnt[k];
        s0 += c.x;
        s1 += c.y;
    }
    unsigned int x0 = s0, x1 = s1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned int u0 = __shfl_up_sync(0xffffffffu, x0, d), u1 = __shfl_up_sync(0xffffffffu, x1, d);
        if (lane >= d) {
            x0 += u0;
            x1 += u1;
        }
    }
    if (lane == 31) {
        ws[0][warp] = x0;
        ws[1][warp] = x1;
    }
    __syncthreads();
    if (warp == 0) {
        const unsigned int v0 = ws[0][lane], v1 = ws[1][lane];
        unsigned int y0 = v0, y1 = v1;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned int u0 = __shfl_up_sync(0xffffffffu, y0, d), u1 = __shfl_up_sync(0xffffffffu, y1, d);
            if (lane >= d) {
                y0 += u0;
                y1 += u1;
            }
        }
        ws[0][lane] = y0 - v0;
        ws[1][lane] = y1 - v1;
    }
    __syncthreads();
    unsigned int o0 = ws[0][warp] + x0 - s0, o1 = ws[1][warp] + x1 - s1;
    for (int k = lo; k < hi; k++) {
        off[k] = make_uint2(o0, o1);
        const uint2 c = cnt[k];
        o0 += c.x;
        o1 += c.y;
    }
}

/* first index of the maximum of ac over [lo, hi) with the running maximum starting at best0 (strict '>': trigger_impl.cc:84-89);
 * -1 when nothing exceeds best0 */
__device__ long long first_record(const float* __restrict__ ac, long long lo, long long hi, float best0, float* best_out)
{
    float best = best0;
    long long idx = -1;
    for (long long i = lo; i < hi; i++) {
        const float v = __ldg(ac + i);
        if (v > best) {
            best = v;
            idx = i;
        }
    }
    *best_out = best;
    return idx;
}

__global__ void __launch_bounds__(64) sync_event_kernel(int nev, const w_det_event* __restrict__ ev, const float* __restrict__ ac,
                                                        const float2* __restrict__ conj, const float2* __restrict__ sig,
                                                        float conj_ac0, float2 conj0, w_det_result* __restrict__ res)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nev) return;
    const w_det_event e = ev[k];
    w_det_result r;
    memset(&r, 0, sizeof(r));

    /* (a) d_conjMultiAvg when the sync block reaches this trigger: the last 0x02 flag in [pos, T), looked for in this
     * event's run and, failing that, in what the previous events left behind */
    float2 cj = conj0;
    for (int j = k; j >= 0; j--) {
        const w_det_event f = ev[j];
        float best;
        const long long idx = first_record(ac, f.rs < 0 ? 0 : f.rs, f.re, f.rs < 0 ? conj_ac0 : 0.0f, &best);
        if (j == k) r.run_max = best;
        if (idx >= 0 && idx >= f.pos) {
            cj = __ldg(conj + idx);
            break;
        }
    }
    r.conj = cj;
    if (!e.do_sync) {
        res[k] = r;
        return;
    }

    /* (b) ltf_autoCorrelation (sync_impl.cc:155-179) */
    const float2* s = sig + e.t;
    cf32 ms = c_make(0.f, 0.f);
    float p1 = 0.f, p2 = 0.f;
    for (int i = 0; i < 64; i++) {
        const float2 a = __ldg(s + i), b = __ldg(s + i + 64);
        ms = c_add(ms, c_mul(c_make(a.x, a.y), c_make(b.x, -b.y)));
        const float aa = c_abs(c_make(a.x, a.y)), bb = c_abs(c_make(b.x, b.y));
        p1 = __fadd_rn(p1, __fmul_rn(aa, aa));
        p2 = __fadd_rn(p2, __fmul_rn(bb, bb));
    }
    float best = 0.f, best_pwr = 0.f;
    int best_i = 0;
    float acv[111];
    for (int i = 0; i < 111; i++) {
        const float v = __fdiv_rn(__fdiv_rn(c_abs(ms), __fsqrt_rn(p1)), __fsqrt_rn(p2));
        acv[i] = v;
        if (i == 0 || best < v) { /* std::max_element: first of the largest */
            best = v;
            best_i = i;
            best_pwr = p1;
        }
        const float2 a = __ldg(s + i), b = __ldg(s + i + 64), c = __ldg(s + i + 128);
        const float aa = c_abs(c_make(a.x, a.y)), bb = c_abs(c_make(b.x, b.y)), cc = c_abs(c_make(c.x, c.y));
        ms = c_sub(ms, c_mul(c_make(a.x, a.y), c_make(b.x, -b.y)));
        p1 = __fsub_rn(p1, __fmul_rn(aa, aa));
        p2 = __fsub_rn(p2, __fmul_rn(bb, bb));
        ms = c_add(ms, c_mul(c_make(b.x, b.y), c_make(c.x, -c.y)));
        p1 = __fadd_rn(p1, __fmul_rn(bb, bb));
        p2 = __fadd_rn(p2, __fmul_rn(cc, cc));
    }
    if (!((double)best > 0.5)) { /* "some miss trigger not higher than 0.5" (sync_impl.cc:99) */
        res[k] = r;
        return;
    }
    const float thr = (float)((double)best * 0.8);
    int li = best_i, ri = best_i;
    for (int j = best_i; j >= 0; j--)
        if (acv[j] < thr) {
            li = j;
            break;
        }
    for (int j = best_i; j < 111; j++)
        if (acv[j] < thr) {
            ri = j;
            break;
        }
    const int mi = (li + ri) / 2;

    /* (c) ltf_cfo (sync_impl.cc:181-196) */
    const float step_stf = __fdiv_rn(atan2f(cj.y, cj.x), 16.0f);
    const float2* m = s + mi;
    cf32 cs = c_make(0.f, 0.f);
    for (int i = 0; i < 64; i++) {
        float s1, c1, s2, c2;
        sincosf(__fmul_rn((float)i, step_stf), &s1, &c1);
        sincosf(__fmul_rn((float)(i + 64), step_stf), &s2, &c2);
        const float2 a = __ldg(m + i), b = __ldg(m + i + 64);
        const cf32 ra = c_mul(c_make(a.x, a.y), c_make(c1, s1));
        const cf32 rb = c_mul(c_make(b.x, b.y), c_make(c2, s2));
        cs = c_add(cs, c_mul(ra, c_conj(rb)));
    }
    const cf32 avg = c_divf(cs, 64.0f);
    const float step_ltf = __fdiv_rn(atan2f(avg.y, avg.x), 64.0f);
    r.valid = 1;
    r.mindex = mi;
    r.rad = __fadd_rn(step_stf, step_ltf);
    const double md = (double)best;
    r.snr = (float)(10.0 * log10(md / (1.0 - md)));
    r.rssi = __fdiv_rn(best_pwr, 64.0f);
    res[k] = r;
}

int w_edges_blocks(unsigned long long n) { return (int)((n + EDGE_BLOCK * EDGE_TILES - 1) / (EDGE_BLOCK * EDGE_TILES)); }

/* counts -> offsets -> ordered lists, all on `st`; counters[0..1] = number of starts / ends (may exceed cap: the caller grows
 * the lists and calls again).  blk: 2 x w_edges_blocks(n) uint2 of scratch. */
void w_launch_run_edges(cudaStream_t st, const float* ac, unsigned long long n, int prev_above0, unsigned int cap,
                        unsigned int* counters, unsigned int* starts, unsigned int* ends, void* blk)
{
    if (n == 0) return;
    const int nb = w_edges_blocks(n);
    uint2* cnt = (uint2*)blk;
    uint2* off = cnt + nb;
    run_edges_kernel<false><<<nb, EDGE_BLOCK, 0, st>>>(ac, n, prev_above0, cnt, nullptr, 0, nullptr, nullptr, counters);
    edges_scan_kernel<<<1, 1024, 0, st>>>(cnt, off, nb);
    run_edges_kernel<true><<<nb, EDGE_BLOCK, 0, st>>>(ac, n, prev_above0, nullptr, off, cap, starts, ends, nullptr);
}

void w_launch_sync_events(cudaStream_t st, int nev, const w_det_event* ev, const float* ac, const float2* conj, const float2* sig,
                          float conj_ac0, float2 conj0, w_det_result* res)
{
    if (nev <= 0) return;
    sync_event_kernel<<<(nev + 63) / 64, 64, 0, st>>>(nev, ev, ac, conj, sig, conj_ac0, conj0, res);
}

/* ------------------------------------------------------------------------------------------------
 * examples/presiso.grc: the two streams the trigger and sync blocks read.
 *   preconj[n] = sum_{k<48} x[n-16-k] * conj(x[n-k])          (delay 16, multiply conjugate, moving average 48)
 *   preac[n]   = |preconj[n]| / sum_{k<64} |x[n-k]|^2         (complex to mag; mag squared, moving average 64; divide)
 * Samples before the start of the buffer count as zero.  The reference has no source for this stage (a GRC hierarchy
 * of stock GNU Radio blocks whose moving averages are running binary32 sums over the whole stream), so results are
 * defined to the 1e-4 tolerance against a binary64 evaluation, not bit for bit.
 * One thread produces PRE_SPAN consecutive outputs: the first by summing its windows, the rest by sliding them, from a
 * shared-memory tile of the block's samples. */
#define PRE_SPAN 8
#define PRE_SHIFT 3 /* log2(PRE_SPAN) */
#define PRE_THREADS 128
#define PRE_TILE (PRE_SPAN * PRE_THREADS)
__global__ void __launch_bounds__(PRE_THREADS) presiso_kernel(const float2* __restrict__ x, unsigned long long n,
                                                              float* __restrict__ ac, float2* __restrict__ cj)
{
    /* one pad element per PRE_SPAN samples: the lanes of a warp are PRE_SPAN samples apart, PRE_SPAN + 1 elements apart in
     * shared memory */
#define T_(i) tile[(i) + ((i) >> PRE_SHIFT)]
    __shared__ float2 tile[PRE_TILE + 80 + (PRE_TILE + 80) / PRE_SPAN + 1];
    const long long base = (long long)blockIdx.x * PRE_TILE;
    for (int i = threadIdx.x; i < PRE_TILE + 80; i += PRE_THREADS) {
        const long long g = base - 80 + i;
        T_(i) = (g >= 0 && g < (long long)n) ? __ldg(x + g) : make_float2(0.f, 0.f);
    }
    __syncthreads();
    const int o = threadIdx.x * PRE_SPAN + 80; /* tile index of this thread's first output sample */
    float sr = 0.f, si = 0.f, pw = 0.f;
    for (int k = 47; k >= 0; k--) { /* oldest first */
        const float2 a = T_(o - 16 - k), b = T_(o - k);
        sr += a.x * b.x + a.y * b.y;
        si += a.y * b.x - a.x * b.y;
    }
    for (int k = 63; k >= 0; k--) {
        const float2 b = T_(o - k);
        pw += b.x * b.x + b.y * b.y;
    }
#pragma unroll
    for (int j = 0; j < PRE_SPAN; j++) {
        const long long g = base + threadIdx.x * PRE_SPAN + j;
        if (g < (long long)n) {
            ac[g] = sqrtf(sr * sr + si * si) / pw;
            cj[g] = make_float2(sr, si);
        }
        /* slide both windows by one sample */
        if (j + 1 < PRE_SPAN) {
            const int m = o + j + 1;
            const float2 na = T_(m - 16), nb = T_(m), oa = T_(m - 64), ob = T_(m - 48), op = T_(m - 64);
            sr += (na.x * nb.x + na.y * nb.y) - (oa.x * ob.x + oa.y * ob.y);
            si += (na.y * nb.x - na.x * nb.y) - (oa.y * ob.x - oa.x * ob.y);
            pw += (nb.x * nb.x + nb.y * nb.y) - (op.x * op.x + op.y * op.y);
        }
    }
}
#undef T_

void w_launch_presiso(cudaStream_t st, const float2* x, unsigned long long n, float* ac, float2* cj)
{
    if (n == 0) return;
    presiso_kernel<<<(unsigned int)((n + PRE_TILE - 1) / PRE_TILE), PRE_THREADS, 0, st>>>(x, n, ac, cj);
}
