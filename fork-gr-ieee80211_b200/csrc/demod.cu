/* demod.cu -- per-symbol OFDM demodulation, 8 lanes per OFDM symbol, 4 symbols per warp.
 *
 * Replaces the DEMOD state of the reference (lib/demod_impl.cc:279-314, lib/demod2_impl.cc:268-330):
 *   fftDemod                      demod_impl.cc:541-547        64-point FFT at symbol start + 8
 *   legacyChanUpdate              demod_impl.cc:507-539        Y/H, pilot phase, reorder
 *   nonLegacyChanUpdate           demod_impl.cc:413-447
 *   htChanUpdate / vhtChanUpdate  demod2_impl.cc:471-630       1-SS and 2x2 zero forcing, joint pilot phase
 *   procSymQamToLlr               cloud80211phy.cc:2090-2148   max-log LLRs, I bits then Q bits
 *   procSymDeintL2 / NL2SS1 / NL2SS2, procSymDepasNL  :2150-2338, :2442-2451  (folded into one index table)
 * The running state of the reference (pilot polarity index, pilot rotation) is a closed form of the
 * symbol number here, so symbols are independent.  A work item is a run of up to W_DEMOD_SPC symbols of one packet, walked by one
 * 8-lane group with the next symbol's samples prefetched; reads: nRx * 64 complex samples per
 * symbol (coalesced 64-byte segments); the symbol's nCBPS LLRs are scattered (deinterleaved) into a
 * staging row in shared memory, which one lane of the group hands to the copy engine of the SM as a single
 * bulk store (cp.async.bulk shared -> global, 192 .. 3328 bytes): no lane spends load / store slots on the
 * copy-out, and the next symbol's FFT and equalisation run while the row drains.
 */
#include "common.cuh"
#include <atomic>

#define DEMOD_THREADS 64   /* one antenna: 8 groups per block, 8 blocks per SM */
#ifndef DEMOD_THREADS2
#define DEMOD_THREADS2 128 /* two antennas: 16 groups per block, 2 blocks per SM (why: at DemodSmem) */
#endif
#ifndef DEMOD_MINBLK
#define DEMOD_MINBLK 8
#endif
#ifndef DEMOD_MINBLK2
#define DEMOD_MINBLK2 2 /* two-antenna variant */
#endif
template <bool MIMO> struct DemodCfg { enum { THREADS = MIMO ? DEMOD_THREADS2 : DEMOD_THREADS, GROUPS = THREADS / 8 }; };
#define DEMOD_SPC W_DEMOD_SPC /* most symbols per work item (one 8-lane group walks them in order) */
#define DEMOD_STAGE 832    /* floats: largest nCBPS (2 streams of 256-QAM) */
#define DEMOD_PAD 4        /* floats after a group's staging row: the slot that bins without data write to.  Rows are as long as the
                              batch's largest symbol (demod_row_floats): an odd number of 16-byte units, so they stay 16-byte aligned
                              for the copy-out and start an odd multiple of 4 banks apart (the deinterleaved positions of the 8 bins
                              a group writes at a time fall on 8 banks that repeat from row to row) */
#define DEMOD_MAPPAD 8     /* map entries after a table, all pointing at that slot */

/* item -> packet: prefix[p] <= w < prefix[p+1].  Batches are mostly homogeneous, so an
 * interpolated guess is tried before the binary search. */
__device__ __forceinline__ int find_packet(const int* __restrict__ prefix, int npkt, int nitems, int w)
{
    int gss = (int)(((long long)w * npkt) / nitems);
    if (__ldg(&prefix[gss]) <= w && w < __ldg(&prefix[gss + 1])) return gss;
    int lo = 0, hi = npkt;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (__ldg(&prefix[mid]) <= w)
            lo = mid;
        else
            hi = mid;
    }
    return lo;
}

/* One bin's row of the scatter table (W80211Tables::binmap): eight 16-bit indices into the group's staging row, bit b of the
 * bin's subcarrier goes to out[idx<b>]; one 16-byte load per bin and lane, straight from global memory (19 KB for all
 * modulations, shared by every block: L1 hits). */
template <int E>
__device__ __forceinline__ uint32_t map_idx(const uint4& m)
{
    const uint32_t w = E < 2 ? m.x : (E < 4 ? m.y : (E < 6 ? m.z : m.w));
    return (E & 1) ? (w >> 16) : (w & 0xffffu);
}

/* max-log LLRs of one axis value (already scaled), written through the scatter table (procSymQamToLlr,
 * cloud80211phy.cc:2090-2148); HALF = bits per axis, E0 = the axis' first table entry: compile-time, so the chain has no
 * branches */
template <int HALF, int E0>
__device__ __forceinline__ void axis_llr(float v, const uint4& m, float* __restrict__ out)
{
    out[map_idx<E0>(m)] = v;
    if (HALF >= 2) {
        float thr = HALF == 2 ? 2.0f : (HALF == 3 ? 4.0f : 8.0f);
        float l = __fsub_rn(thr, fabsf(v));
        out[map_idx<E0 + 1>(m)] = l;
        if (HALF >= 3) {
            thr = __fmul_rn(thr, 0.5f);
            l = __fsub_rn(thr, fabsf(l));
            out[map_idx<E0 + 2>(m)] = l;
            if (HALF >= 4) {
                l = __fsub_rn(2.0f, fabsf(l));
                out[map_idx<E0 + 3>(m)] = l;
            }
        }
    }
}

template <int NB>
__device__ __forceinline__ void qam_llr(cf32 q, const uint4& m, float* __restrict__ out)
{
    if (NB == 1) {
        out[map_idx<0>(m)] = q.x;
        return;
    }
    /* reference constants, cloud80211phy.cc:2105,2114,2125,2138 */
    const float sc = NB == 2 ? 1.4142135623730951f : (NB == 4 ? 3.1622776601683795f : (NB == 6 ? 6.48074069840786f : 13.038404810405298f));
    q = c_scale(q, sc);
    axis_llr<(NB >> 1), 0>(q.x, m, out);
    axis_llr<(NB >> 1), (NB >> 1)>(q.y, m, out);
}

/* The LLRs of one symbol: q = eq * conj(pilot sum) / |pilot sum| per data subcarrier, demapped and scattered into the
 * group's staging block in coded-bit order.  No branch per bin: the table rows of a bin without data point at the dump slot
 * after the symbol's LLRs.  q = a / pa goes through y = RN(1/pa) (f_div_known); unsafe divisors and tiny quotients
 * (where the FMA residual could leave the normal range) are found per symbol, not per bin, and the whole symbol is then
 * redone with the plain division. */
template <int NB, bool MIMO, bool SLOW>
__device__ __forceinline__ bool symbol_llrs_pass(const cf32 (&e1)[8], const cf32 (&e2)[8], const int (&jj)[8], bool two, cf32 ps, float pa,
                                                 float rpa, const uint4* __restrict__ map1, const uint4* __restrict__ map2,
                                                 float* __restrict__ stg)
{
    bool bad = false;
#pragma unroll
    for (int k2 = 0; k2 < 8; k2++) {
        const int j = jj[k2];
        {
            const uint4 m = __ldg(map1 + 8 * k2);
            const cf32 a = c_mul(e1[k2], ps);
            const cf32 q = SLOW ? c_divf(a, pa) : c_make(f_div_known(a.x, pa, rpa), f_div_known(a.y, pa, rpa));
            if (!SLOW) bad = bad || (j >= 0 && !(fabsf(q.x) >= 1e-30f && fabsf(q.y) >= 1e-30f));
            qam_llr<NB>(q, m, stg);
        }
        if (MIMO && two) {
            const uint4 m = __ldg(map2 + 8 * k2);
            const cf32 a = c_mul(e2[k2], ps);
            const cf32 q = SLOW ? c_divf(a, pa) : c_make(f_div_known(a.x, pa, rpa), f_div_known(a.y, pa, rpa));
            if (!SLOW) bad = bad || (j >= 0 && !(fabsf(q.x) >= 1e-30f && fabsf(q.y) >= 1e-30f));
            qam_llr<NB>(q, m, stg);
        }
    }
    return bad;
}

template <int NB, bool MIMO>
__device__ __forceinline__ void symbol_llrs(const cf32 (&e1)[8], const cf32 (&e2)[8], const int (&jj)[8], bool two, cf32 ps, float pa,
                                            const uint4* __restrict__ map1, const uint4* __restrict__ map2, float* __restrict__ stg)
{
    const bool slow_pa = f_recip_unsafe(pa) || !(pa > 1e-30f) || !(pa < 1e30f);
    const float rpa = __frcp_rn(pa);
    bool redo = slow_pa;
    if (!slow_pa) redo = symbol_llrs_pass<NB, MIMO, false>(e1, e2, jj, two, ps, pa, rpa, map1, map2, stg);
    if (redo) symbol_llrs_pass<NB, MIMO, true>(e1, e2, jj, two, ps, pa, rpa, map1, map2, stg);
}

__device__ __forceinline__ float shfl_f(float v, int src) { return __shfl_sync(0xffffffffu, v, src); }

__device__ __forceinline__ void load_sym(const float2* __restrict__ src, int g, bool valid, float2* v)
{
#pragma unroll
    for (int n1 = 0; n1 < 8; n1++) v[n1] = valid ? __ldcs(&src[8 * n1 + g]) : make_float2(0.f, 0.f); /* streamed once: evict-first, the tables and channel blocks stay in L1 */
}

/* FFT of samples already in registers (same arithmetic as fft64_lane8); the lane's twiddles W64^(g*k1) come from the
 * block's shared copy instead of sixteen registers per thread */
__device__ __forceinline__ void fft64_regs(const float2* v, int g, float2* xch, float* re, float* im, const float2* __restrict__ tw /* [g][9] */)
{
#pragma unroll
    for (int n1 = 0; n1 < 8; n1++) {
        re[n1] = v[n1].x;
        im[n1] = v[n1].y;
    }
    c8_fft8(re, im);
    if (g) {
#pragma unroll
        for (int k1 = 1; k1 < 8; k1++) {
            const float2 w = tw[g * 9 + k1];
            float ar = re[k1], ai = im[k1];
            re[k1] = C8_SUB(C8_MUL(ar, w.x), C8_MUL(ai, w.y));
            im[k1] = C8_ADD(C8_MUL(ar, w.y), C8_MUL(ai, w.x));
        }
    }
#pragma unroll
    for (int k1 = 0; k1 < 8; k1++) xch[g * 9 + k1] = make_float2(re[k1], im[k1]);
    __syncwarp();
#pragma unroll
    for (int n2 = 0; n2 < 8; n2++) {
        float2 t = xch[n2 * 9 + g];
        re[n2] = t.x;
        im[n2] = t.y;
    }
    __syncwarp();
    c8_fft8(re, im);
}

/* dynamic shared memory of one block (the two-antenna variant needs more than the 48 KB static limit) */
template <bool MIMO>
struct DemodSmem {
    enum { GROUPS = DemodCfg<MIMO>::GROUPS };
    double eqr[MIMO ? 1 : GROUPS][MIMO ? 1 : 64]; /* per bin: RN(1 / |H|^2) in binary64.  Not kept by the two-antenna variant (its
                                                     one-stream packets recompute it) */
    alignas(16) float2 xch[GROUPS][W_FFT_XCH_FLOAT2]; /* FFT exchange area of the group (the staging row may still be draining when
                                                         the next symbol's FFT runs) */
    float2 tw[8 * 9];                                 /* W64^(g*k1) at [g*9 + k1] */
    alignas(16) float stage[4];                       /* GROUPS rows of row_floats floats follow: one symbol's LLRs per group,
                                                         deinterleaved, the source of the bulk store */
};
/* Shared memory decides how much L1 the SM keeps (the carve-out is the smallest step that holds the resident blocks), and the
 * two-antenna variant lives on L1 (channel matrices, table rows: profiles/experiments/r2_demod.txt), so the rows are sized by
 * the batch, not by the largest symbol of the standard: two blocks of 16 groups of C5 (nCBPS 624) fit the 100 KB step, which
 * four blocks of 8 groups miss by the second kilobyte the driver reserves per block. */
__host__ __device__ inline int demod_row_floats(int max_ncbps)
{
    int r = ((max_ncbps + 3) & ~3) + DEMOD_PAD;
    if (((r >> 2) & 1) == 0) r += 4;
    return r;
}
template <bool MIMO> static size_t demod_smem_bytes(int max_ncbps)
{
    return offsetof(DemodSmem<MIMO>, stage) + (size_t)DemodCfg<MIMO>::GROUPS * demod_row_floats(max_ncbps) * sizeof(float);
}
static_assert(offsetof(W80211Tables, binmap) % 16 == 0, "binmap rows are read as uint4");

/* shared -> global bulk copy by the SM's copy engine, tracked by this thread's bulk async-group (PTX cp.async.bulk) */
__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"((uint32_t)__cvta_generic_to_shared(ssrc)), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
/* until the copy engine has READ the source of every bulk store this thread has issued (the row may be overwritten then) */
__device__ __forceinline__ void bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

template <bool MIMO> /* MIMO = false: one antenna, no two-stream code or registers at all */
__global__ void __launch_bounds__(DemodCfg<MIMO>::THREADS, MIMO ? DEMOD_MINBLK2 : DEMOD_MINBLK) demod_kernel(const W80211Tables* __restrict__ T, int nitems, int npkt, int nant, int row_floats,
                                                              const float2* __restrict__ iq0, const float2* __restrict__ iq1,
                                                              const w80211_pkt_in* __restrict__ pkin,
                                                              const w80211_desc* __restrict__ desc, const int* __restrict__ item_prefix,
                                                              const float2* __restrict__ chan, int chan_stride, float* __restrict__ llr)
{
    extern __shared__ __align__(16) unsigned char demod_smem_raw[];
    DemodSmem<MIMO>& S = *reinterpret_cast<DemodSmem<MIMO>*>(demod_smem_raw);
    auto& eqr = S.eqr;
    const int tid = threadIdx.x;
    const int lane = tid & 31, g = tid & 7, base = lane & ~7, grp = tid >> 3;
    const int w = blockIdx.x * DemodCfg<MIMO>::GROUPS + grp;
    const bool valid = w < nitems;
    if (tid < 64) S.tw[(tid >> 3) * 9 + (tid & 7)] = make_float2(C8_W64_RE[((tid >> 3) * (tid & 7)) & 63], C8_W64_IM[((tid >> 3) * (tid & 7)) & 63]);
    __syncthreads();
    const float2* tw = S.tw;

    int p = 0, n0 = 0, nsy = 0;
    int fmt = 0, nss = 1, nb = 1, ncbps = 48, nsymsamp = 80;
    long long llr_off = 0;
    const float2* src0 = iq0;
    const float2* src1 = iq0;
    if (valid) {
        p = find_packet(item_prefix, npkt, nitems, w);
        const w80211_desc* d = &desc[p];
        /* the packet's symbols are dealt out evenly over its items (105 symbols, 3 items: 35 each, not 36 + 36 + 33), so
         * the four groups of a warp, which loop together, finish together */
        const int first = __ldg(&item_prefix[p]), nit = __ldg(&item_prefix[p + 1]) - first;
        const int per = (d->nsym + nit - 1) / nit;
        n0 = (w - first) * per;
        nsy = max(0, min(per, d->nsym - n0));
        fmt = d->format;
        nss = d->nss;
        nb = d->nbpscs;
        ncbps = d->ncbps;
        nsymsamp = d->nsymsamp;
        llr_off = d->llr_off;
        size_t off = (size_t)pkin[p].iq_offset + d->data_off + (size_t)n0 * nsymsamp + 8;
        src0 = iq0 + off;
        src1 = nant == 2 ? iq1 + off : src0;
    }
    const bool two = MIMO && valid && nss == 2;
    const bool any2 = MIMO && __any_sync(0xffffffffu, two);
    const int nloop = __reduce_max_sync(0xffffffffu, nsy);
    float* stg = S.stage + (size_t)grp * row_floats;
    float2* myx = S.xch[grp];

    /* per-item constants: channel of the lane's eight bins (one stream), tables, pilot patterns */
    const float2* H = chan + (size_t)p * chan_stride;
    const int8_t* binj = fmt == W80211_F_L ? T->bin_j_l : T->bin_j_nl;
    int jj[8];
    cf32 hh[8];
    bool slow_div = false; /* a divisor with an all-ones significand: take the plain division (never seen in practice) */
#pragma unroll
    for (int k2 = 0; k2 < 8; k2++) {
        int i = g + 8 * k2;
        jj[k2] = valid ? binj[i] : -1;
        bool used = valid && (jj[k2] >= 0 || i == 7 || i == 21 || i == 43 || i == 57);
        if (!used) jj[k2] = -2; /* -1: pilot (equalised, no LLR), -2: unused bin */
        hh[k2] = c_make(1.f, 0.f);
        if (used && !two) {
            float2 h = __ldg(&H[i]);
            hh[k2] = c_make(h.x, h.y);
        }
        /* the divisor of Y/H is the same for every symbol of the packet: |H|^2 and RN(1/|H|^2) once per item */
        const double hc = hh[k2].x, hd = hh[k2].y;
        const double denom = __dadd_rn(__dmul_rn(hc, hc), __dmul_rn(hd, hd));
        slow_div = slow_div || d_recip_unsafe(denom);
        if (!MIMO) eqr[grp][i] = __drcp_rn(denom);
    }
    const int z = fmt == W80211_F_L ? 1 : (fmt == W80211_F_HT ? 3 : 4);
    const bool ht2 = two && fmt == W80211_F_HT;
    const int nbi = nb == 1 ? 0 : (nb == 2 ? 1 : (nb == 4 ? 2 : (nb == 6 ? 3 : 4)));
    const int tid1 = fmt == W80211_F_L ? nbi : (two ? 9 + nbi : 4 + nbi);
    const uint4* map1 = reinterpret_cast<const uint4*>(T->binmap[tid1]) + g;       /* row of bin g + 8 k2 at [8 k2] */
    const uint4* map2 = reinterpret_cast<const uint4*>(T->binmap[14 + nbi]) + g;
    cf32 R[2][4];
#pragma unroll
    for (int s2 = 0; s2 < 2; s2++)
#pragma unroll
        for (int k = 0; k < 4; k++) {
            R[s2][k] = c_make(1.f, 0.f);
            if (two) {
                float2 v = __ldg(&H[512 + 4 * s2 + k]);
                R[s2][k] = c_make(v.x, v.y);
            }
        }

    float2 cur0[8], cur1[8];
    load_sym(src0, g, nsy > 0, cur0);
    if (any2) load_sym(src1, g, two && nsy > 0, cur1);
    for (int it = 0; it < nloop; it++) {
        const bool act = it < nsy;
        const int n = n0 + it;
        float2 nx0[8], nx1[8];
        const bool more = it + 1 < nsy;
        load_sym(src0 + (size_t)(it + 1) * nsymsamp, g, more, nx0);
        if (any2) load_sym(src1 + (size_t)(it + 1) * nsymsamp, g, two && more, nx1);

        float re[8], im[8], re2[8], im2[8];
        fft64_regs(cur0, g, myx, re, im, tw);
        if (any2) fft64_regs(cur1, g, myx, re2, im2, tw);

        cf32 e1[8], e2[8];
#pragma unroll
        for (int k2 = 0; k2 < 8; k2++) e2[k2] = c_make(0.f, 0.f);
        if (!MIMO || !two) {
#pragma unroll
            for (int k2 = 0; k2 < 8; k2++) {
                const int i = g + 8 * k2;
                /* every bin, used or not (unused bins divide by 1): no branch in the one-stream path.
                 * keep the conversions of H and |H|^2 inside the loop: hoisted, they cost 48 registers */
                asm volatile("" : "+f"(hh[k2].x), "+f"(hh[k2].y));
                const double hc = hh[k2].x, hd = hh[k2].y;
                const double denom = __dadd_rn(__dmul_rn(hc, hc), __dmul_rn(hd, hd));
                e1[k2] = slow_div ? c_div(c_make(re[k2], im[k2]), hh[k2]) : c_div_known(c_make(re[k2], im[k2]), hh[k2], denom, MIMO ? __drcp_rn(denom) : eqr[MIMO ? 0 : grp][MIMO ? 0 : i]);
            }
        } else {
            /* two streams: H^H y, then the inverse of H^H H (demod2_impl.cc:471-630), every bin and every lane alike -- no
             * condition between the loads and the loop head, so they are issued bins ahead of their use (the results of unused
             * bins and of lanes past their item's end go nowhere: their table rows point at the dump slot, nothing is stored
             * for an inactive lane) */
#pragma unroll
            for (int k2 = 0; k2 < 8; k2++) {
                const int i = g + 8 * k2;
                const float4* H4 = reinterpret_cast<const float4*>(H + i * 4);
                const float4* I4 = reinterpret_cast<const float4*>(H + 256 + i * 4);
                const float4 ha = __ldg(H4), hb = __ldg(H4 + 1), ia = __ldg(I4), ib = __ldg(I4 + 1);
                const cf32 y1 = c_make(re[k2], im[k2]), y2 = c_make(re2[k2], im2[k2]);
                const cf32 t1 = c_add(c_mul(y1, c_make(ha.x, -ha.y)), c_mul(y2, c_make(ha.z, -ha.w)));
                const cf32 t2 = c_add(c_mul(y1, c_make(hb.x, -hb.y)), c_mul(y2, c_make(hb.z, -hb.w)));
                e1[k2] = c_add(c_mul(t1, c_make(ia.x, ia.y)), c_mul(t2, c_make(ib.x, ib.y)));
                e2[k2] = c_add(c_mul(t1, c_make(ia.z, ia.w)), c_mul(t2, c_make(ib.z, ib.w)));
            }
        }
        /* the four pilots of each stream, to every lane of the group: bins 7, 21, 43, 57 */
        cf32 a7 = c_make(shfl_f(e1[0].x, base + 7), shfl_f(e1[0].y, base + 7));
        cf32 a21 = c_make(shfl_f(e1[2].x, base + 5), shfl_f(e1[2].y, base + 5));
        cf32 a43 = c_make(shfl_f(e1[5].x, base + 3), shfl_f(e1[5].y, base + 3));
        cf32 a57 = c_make(shfl_f(e1[7].x, base + 1), shfl_f(e1[7].y, base + 1));
        cf32 b7 = c_make(0.f, 0.f), b21 = b7, b43 = b7, b57 = b7;
        if (any2) {
            b7 = c_make(shfl_f(e2[0].x, base + 7), shfl_f(e2[0].y, base + 7));
            b21 = c_make(shfl_f(e2[2].x, base + 5), shfl_f(e2[2].y, base + 5));
            b43 = c_make(shfl_f(e2[5].x, base + 3), shfl_f(e2[5].y, base + 3));
            b57 = c_make(shfl_f(e2[7].x, base + 1), shfl_f(e2[7].y, base + 1));
        }
        cf32 ps = c_make(1.f, 0.f);
        float pa = 1.f;
        if (act) {
            /* pilot polarity p_{z+n} and the per-symbol rotation of the base pattern (pilotShift) */
            const float P = T->pilot_p[(z + n) % 127];
            const int rot = fmt == W80211_F_L ? 0 : (n & 3);
            float pl[4], pl2[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                int idx = (k + rot) & 3;
                float v1 = ht2 ? (idx < 2 ? 1.0f : -1.0f) : (idx == 3 ? -1.0f : 1.0f);  /* PILOT_HT_2_1 : PILOT_L/HT_1/VHT */
                float v2 = ht2 ? ((idx == 0 || idx == 3) ? 1.0f : -1.0f) : v1;          /* PILOT_HT_2_2 : same as stream 1 */
                pl[k] = v1;
                pl2[k] = v2;
            }
            cf32 sum;
            if (!two) {
                sum = c_scale(c_scale(a7, pl[2]), P);
                sum = c_add(sum, c_scale(c_scale(a21, pl[3]), P));
                sum = c_add(sum, c_scale(c_scale(a43, pl[0]), P));
                sum = c_add(sum, c_scale(c_scale(a57, pl[1]), P));
            } else {
                sum = c_mul(c_scale(c_scale(a7, pl[2]), P), R[0][2]);
                sum = c_add(sum, c_mul(c_scale(c_scale(a21, pl[3]), P), R[0][3]));
                sum = c_add(sum, c_mul(c_scale(c_scale(a43, pl[0]), P), R[0][0]));
                sum = c_add(sum, c_mul(c_scale(c_scale(a57, pl[1]), P), R[0][1]));
                sum = c_add(sum, c_mul(c_scale(c_scale(b7, pl2[2]), P), R[1][2]));
                sum = c_add(sum, c_mul(c_scale(c_scale(b21, pl2[3]), P), R[1][3]));
                sum = c_add(sum, c_mul(c_scale(c_scale(b43, pl2[0]), P), R[1][0]));
                sum = c_add(sum, c_mul(c_scale(c_scale(b57, pl2[1]), P), R[1][1]));
            }
            ps = c_conj(sum);
            pa = c_abs(ps);
        }
        /* the previous symbol's row must have left shared memory before this symbol's LLRs go into it */
        if (g == 0) bulk_store_wait_read();
        __syncwarp();
        if (act) {
            /* LLRs through the combined deinterleave / de-parse table into the shared staging block; one instance
             * per modulation so the demapping chain is straight-line code */
            switch (nb) {
            case 1: symbol_llrs<1, MIMO>(e1, e2, jj, two, ps, pa, map1, map2, stg); break;
            case 2: symbol_llrs<2, MIMO>(e1, e2, jj, two, ps, pa, map1, map2, stg); break;
            case 4: symbol_llrs<4, MIMO>(e1, e2, jj, two, ps, pa, map1, map2, stg); break;
            case 6: symbol_llrs<6, MIMO>(e1, e2, jj, two, ps, pa, map1, map2, stg); break;
            default: symbol_llrs<8, MIMO>(e1, e2, jj, two, ps, pa, map1, map2, stg); break;
            }
            /* every lane's writes become visible to the copy engine (async proxy), then one lane hands it the row: the symbol's
             * nCBPS floats, 16-byte aligned on both sides (llr_off and nCBPS are multiples of 4 floats) */
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
        __syncwarp();
        if (act && g == 0) bulk_store(llr + llr_off + (size_t)n * ncbps, stg, (uint32_t)ncbps * 4u);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            cur0[k] = nx0[k];
            cur1[k] = nx1[k];
        }
    }
    if (g == 0) bulk_store_wait_read(); /* the block's shared memory must outlive the copy engine's reads */
}

/* Function attributes are per device: w80211_create() calls this once for the context's device (the current one). */
int w_demod_init_device()
{
    cudaError_t e = cudaFuncSetAttribute(demod_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)demod_smem_bytes<true>(DEMOD_STAGE));
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(demod_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)demod_smem_bytes<false>(DEMOD_STAGE / 2));
    return e == cudaSuccess ? 0 : -1;
}

/* Ask for the smallest shared-memory step that holds `blocks` resident blocks (the rest of the 256 KB stays L1); left alone
 * the driver takes the 132 KB step for two 50 KB blocks that fit the 100 KB one.  The attribute belongs to the function on a
 * device: it is set again only when the batch's shape changes the answer (two host threads with different shapes may
 * overwrite each other's choice -- that costs L1, never correctness). */
template <bool MIMO> static void demod_prefer_carveout(size_t smem_per_block, int blocks)
{
    static std::atomic<int> last[64];
    const size_t need = (smem_per_block + 1024) * (size_t)blocks; /* the driver reserves 1 KB per block */
    static const int steps_kb[] = {8, 16, 32, 64, 100, 132, 164, 196, 228};
    int kb = 228;
    for (int c : steps_kb)
        if ((size_t)c * 1024 >= need) {
            kb = c;
            break;
        }
    const int pct = kb * 100 / 228; /* rounded down: the driver rounds the hint up to the next step */
    int dev = 0;
    cudaGetDevice(&dev);
    if (last[dev & 63].exchange(pct + 1) == pct + 1) return;
    cudaFuncSetAttribute(demod_kernel<MIMO>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
}

/* max_ncbps: the largest nCBPS among the batch's packets (sizes the staging rows; anything from 48 to 832) */
void w_launch_demod(cudaStream_t st, const W80211Tables* tab, int nitems, int npkt, int nant, const float2* iq0,
                    const float2* iq1, const w80211_pkt_in* pk, const w80211_desc* desc, const int* item_prefix,
                    const float2* chan, int chan_stride, float* llr, int max_ncbps)
{
    if (nitems <= 0) return;
    if (nant == 2) {
        max_ncbps = max_ncbps < 48 ? 48 : (max_ncbps > DEMOD_STAGE ? DEMOD_STAGE : max_ncbps);
        const int G = DemodCfg<true>::GROUPS;
        demod_prefer_carveout<true>(demod_smem_bytes<true>(max_ncbps), DEMOD_MINBLK2);
        demod_kernel<true><<<(nitems + G - 1) / G, DemodCfg<true>::THREADS, demod_smem_bytes<true>(max_ncbps), st>>>(
            tab, nitems, npkt, nant, demod_row_floats(max_ncbps), iq0, iq1, pk, desc, item_prefix, chan, chan_stride, llr);
    } else {
        max_ncbps = max_ncbps < 48 ? 48 : (max_ncbps > DEMOD_STAGE / 2 ? DEMOD_STAGE / 2 : max_ncbps);
        const int G = DemodCfg<false>::GROUPS;
        demod_prefer_carveout<false>(demod_smem_bytes<false>(max_ncbps), DEMOD_MINBLK);
        demod_kernel<false><<<(nitems + G - 1) / G, DemodCfg<false>::THREADS, demod_smem_bytes<false>(max_ncbps), st>>>(
            tab, nitems, npkt, nant, demod_row_floats(max_ncbps), iq0, iq1, pk, desc, item_prefix, chan, chan_stride, llr);
    }
}
