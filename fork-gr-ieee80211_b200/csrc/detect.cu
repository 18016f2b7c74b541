/* detect.cu -- packet detection on the GPU: the data-parallel parts of the reference's trigger
 * block (lib/trigger_impl.cc:60-117) and sync block (lib/sync_impl.cc:61-196).
 *
 *   run_edges_kernel   one thread per sample: where runs of preac > 0.3 begin and end (the only thing the trigger's
 *                      plateau counter, its running maximum and its flags depend on).  The few edges go to two lists.
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
#include <cub/device/device_radix_sort.cuh>
#include "common.cuh"

__device__ __forceinline__ bool above_thr(float v) { return v > 0.3f; }

__global__ void __launch_bounds__(256) run_edges_kernel(const float* __restrict__ ac, unsigned long long n, int prev_above0,
                                                        unsigned int cap, unsigned int* __restrict__ counters,
                                                        unsigned int* __restrict__ starts, unsigned int* __restrict__ ends)
{
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bool a = above_thr(__ldg(ac + i));
    const bool b = i ? above_thr(__ldg(ac + i - 1)) : (prev_above0 != 0);
    if (a && !b) {
        unsigned int k = atomicAdd(&counters[0], 1u);
        if (k < cap) starts[k] = (unsigned int)i;
    } else if (!a && b) {
        unsigned int k = atomicAdd(&counters[1], 1u);
        if (k < cap) ends[k] = (unsigned int)i;
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

void w_launch_run_edges(cudaStream_t st, const float* ac, unsigned long long n, int prev_above0, unsigned int cap,
                        unsigned int* counters, unsigned int* starts, unsigned int* ends)
{
    if (n == 0) return;
    const unsigned long long blocks = (n + 255) / 256;
    run_edges_kernel<<<(unsigned int)blocks, 256, 0, st>>>(ac, n, prev_above0, cap, counters, starts, ends);
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

/* the edge lists come out of the atomics in arbitrary order: radix sort them on the device (a few 10^4 keys) */
size_t w_sort_u32_temp_bytes(int n)
{
    size_t b = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, b, (const unsigned int*)nullptr, (unsigned int*)nullptr, n);
    return b;
}
void w_sort_u32(cudaStream_t st, void* temp, size_t temp_bytes, const unsigned int* in, unsigned int* out, int n)
{
    if (n > 0) cub::DeviceRadixSort::SortKeys(temp, temp_bytes, in, out, n, 0, 32, st);
}
