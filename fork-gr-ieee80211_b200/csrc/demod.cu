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
 * symbol number here, so symbols are independent.  Reads: nRx * 64 complex samples per symbol
 * (coalesced 64-byte segments); writes: nCBPS floats per symbol, scattered inside the symbol's block.
 */
#include "common.cuh"

#define DEMOD_THREADS 128

__device__ __forceinline__ int find_packet(const int* __restrict__ prefix, int npkt, int w)
{
    int lo = 0, hi = npkt; /* prefix[lo] <= w < prefix[hi] */
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (__ldg(&prefix[mid]) <= w)
            lo = mid;
        else
            hi = mid;
    }
    return lo;
}

/* max-log LLRs of one axis value (already scaled), written through the deinterleave table */
__device__ __forceinline__ void axis_llr(float v, int half, const uint16_t* __restrict__ map, int idx, float* __restrict__ out)
{
    out[map[idx]] = v;
    if (half >= 2) {
        float thr = half == 2 ? 2.0f : (half == 3 ? 4.0f : 8.0f);
        float l = __fsub_rn(thr, fabsf(v));
        out[map[idx + 1]] = l;
        if (half >= 3) {
            thr = __fmul_rn(thr, 0.5f);
            l = __fsub_rn(thr, fabsf(l));
            out[map[idx + 2]] = l;
            if (half >= 4) {
                l = __fsub_rn(2.0f, fabsf(l));
                out[map[idx + 3]] = l;
            }
        }
    }
}

__device__ __forceinline__ void qam_llr(cf32 q, int nb, const uint16_t* __restrict__ map, int j, float* __restrict__ out)
{
    if (nb == 1) {
        out[map[j]] = q.x;
        return;
    }
    /* reference constants, cloud80211phy.cc:2105,2114,2125,2138 */
    float sc = nb == 2 ? 1.4142135623730951f : (nb == 4 ? 3.1622776601683795f : (nb == 6 ? 6.48074069840786f : 13.038404810405298f));
    q = c_scale(q, sc);
    int half = nb >> 1;
    axis_llr(q.x, half, map, j * nb, out);
    axis_llr(q.y, half, map, j * nb + half, out);
}

__device__ __forceinline__ float shfl_f(float v, int src) { return __shfl_sync(0xffffffffu, v, src); }

__global__ void __launch_bounds__(DEMOD_THREADS) demod_kernel(const W80211Tables* __restrict__ T, int nsymtot, int npkt, int nant,
                                                              const float2* __restrict__ iq0, const float2* __restrict__ iq1,
                                                              const w80211_pkt_in* __restrict__ pkin,
                                                              const w80211_desc* __restrict__ desc, const int* __restrict__ sym_prefix,
                                                              const float2* __restrict__ chan, int chan_stride, float* __restrict__ llr)
{
    __shared__ float2 xch[DEMOD_THREADS / 8][W_FFT_XCH_FLOAT2];
    const int tid = threadIdx.x;
    const int lane = tid & 31, g = tid & 7, base = lane & ~7;
    const int w = (blockIdx.x * DEMOD_THREADS + tid) >> 3;
    const bool valid = w < nsymtot;
    float wr[8], wi[8];
    fft64_twiddles(g, wr, wi);

    int p = 0, n = 0;
    int fmt = 0, nss = 1, nb = 1, ncbps = 48, nsymsamp = 80;
    long long llr_off = 0;
    const float2* src0 = iq0;
    const float2* src1 = iq0;
    if (valid) {
        p = find_packet(sym_prefix, npkt, w);
        n = w - __ldg(&sym_prefix[p]);
        const w80211_desc* d = &desc[p];
        fmt = d->format;
        nss = d->nss;
        nb = d->nbpscs;
        ncbps = d->ncbps;
        nsymsamp = d->nsymsamp;
        llr_off = d->llr_off;
        size_t off = (size_t)pkin[p].iq_offset + d->data_off + (size_t)n * nsymsamp + 8;
        src0 = iq0 + off;
        src1 = nant == 2 ? iq1 + off : src0;
    }
    float2* myx = xch[tid >> 3];
    float re[8], im[8];
    fft64_lane8(src0, g, myx, re, im, wr, wi, valid);
    const bool two = valid && nss == 2;
    float re2[8], im2[8];
    if (__any_sync(0xffffffffu, two)) {
        fft64_lane8(src1, g, myx, re2, im2, wr, wi, two);
    }

    /* equalise: lane g owns bins i = g + 8*k2 */
    const float2* H = chan + (size_t)p * chan_stride;
    cf32 e1[8], e2[8];
    const int8_t* binj = fmt == W80211_F_L ? T->bin_j_l : T->bin_j_nl;
    int jj[8];
#pragma unroll
    for (int k2 = 0; k2 < 8; k2++) {
        int i = g + 8 * k2;
        e1[k2] = c_make(0.f, 0.f);
        e2[k2] = c_make(0.f, 0.f);
        jj[k2] = valid ? binj[i] : -1;
        /* used bins: data bins plus the four pilots */
        bool used = valid && (jj[k2] >= 0 || i == 7 || i == 21 || i == 43 || i == 57);
        if (used) {
            if (!two) {
                float2 h = __ldg(&H[nss == 2 ? i * 4 : i]);
                e1[k2] = c_div(c_make(re[k2], im[k2]), c_make(h.x, h.y));
            } else {
                const float4* H4 = reinterpret_cast<const float4*>(H + i * 4);
                const float4* I4 = reinterpret_cast<const float4*>(H + 256 + i * 4);
                float4 ha = __ldg(H4), hb = __ldg(H4 + 1), ia = __ldg(I4), ib = __ldg(I4 + 1);
                cf32 y1 = c_make(re[k2], im[k2]), y2 = c_make(re2[k2], im2[k2]);
                cf32 t1 = c_add(c_mul(y1, c_make(ha.x, -ha.y)), c_mul(y2, c_make(ha.z, -ha.w)));
                cf32 t2 = c_add(c_mul(y1, c_make(hb.x, -hb.y)), c_mul(y2, c_make(hb.z, -hb.w)));
                e1[k2] = c_add(c_mul(t1, c_make(ia.x, ia.y)), c_mul(t2, c_make(ib.x, ib.y)));
                e2[k2] = c_add(c_mul(t1, c_make(ia.z, ia.w)), c_mul(t2, c_make(ib.z, ib.w)));
            }
        }
    }
    /* the four pilots of each stream, to every lane of the group: bins 7, 21, 43, 57 */
    cf32 a7 = c_make(shfl_f(e1[0].x, base + 7), shfl_f(e1[0].y, base + 7));
    cf32 a21 = c_make(shfl_f(e1[2].x, base + 5), shfl_f(e1[2].y, base + 5));
    cf32 a43 = c_make(shfl_f(e1[5].x, base + 3), shfl_f(e1[5].y, base + 3));
    cf32 a57 = c_make(shfl_f(e1[7].x, base + 1), shfl_f(e1[7].y, base + 1));
    cf32 b7 = c_make(0.f, 0.f), b21 = b7, b43 = b7, b57 = b7;
    if (__any_sync(0xffffffffu, two)) {
        b7 = c_make(shfl_f(e2[0].x, base + 7), shfl_f(e2[0].y, base + 7));
        b21 = c_make(shfl_f(e2[2].x, base + 5), shfl_f(e2[2].y, base + 5));
        b43 = c_make(shfl_f(e2[5].x, base + 3), shfl_f(e2[5].y, base + 3));
        b57 = c_make(shfl_f(e2[7].x, base + 1), shfl_f(e2[7].y, base + 1));
    }
    if (!valid) return;

    /* pilot polarity p_{z+n} and the per-symbol rotation of the base pattern (pilotShift) */
    const int z = fmt == W80211_F_L ? 1 : (fmt == W80211_F_HT ? 3 : 4);
    const float P = T->pilot_p[(z + n) % 127];
    const int rot = fmt == W80211_F_L ? 0 : (n & 3);
    /* base patterns in frequency order -21,-7,+7,+21 (cloud80211phy.cc:170-174) */
    const bool ht2 = two && fmt == W80211_F_HT;
    float pl[4], pl2[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        int idx = (k + rot) & 3;
        float v1 = ht2 ? (idx < 2 ? 1.0f : -1.0f) : (idx == 3 ? -1.0f : 1.0f);       /* PILOT_HT_2_1 : PILOT_L/HT_1/VHT */
        float v2 = ht2 ? ((idx == 0 || idx == 3) ? 1.0f : -1.0f) : v1;               /* PILOT_HT_2_2 : same as stream 1 */
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
        const float2* PR = H + 512;
        auto R = [&](int s, int k) { float2 v = __ldg(&PR[4 * s + k]); return c_make(v.x, v.y); };
        sum = c_mul(c_scale(c_scale(a7, pl[2]), P), R(0, 2));
        sum = c_add(sum, c_mul(c_scale(c_scale(a21, pl[3]), P), R(0, 3)));
        sum = c_add(sum, c_mul(c_scale(c_scale(a43, pl[0]), P), R(0, 0)));
        sum = c_add(sum, c_mul(c_scale(c_scale(a57, pl[1]), P), R(0, 1)));
        sum = c_add(sum, c_mul(c_scale(c_scale(b7, pl2[2]), P), R(1, 2)));
        sum = c_add(sum, c_mul(c_scale(c_scale(b21, pl2[3]), P), R(1, 3)));
        sum = c_add(sum, c_mul(c_scale(c_scale(b43, pl2[0]), P), R(1, 0)));
        sum = c_add(sum, c_mul(c_scale(c_scale(b57, pl2[1]), P), R(1, 1)));
    }
    const cf32 ps = c_conj(sum);
    const float pa = c_abs(ps);

    /* LLRs through the combined deinterleave / de-parse table, into the symbol's nCBPS block */
    const int nbi = nb == 1 ? 0 : (nb == 2 ? 1 : (nb == 4 ? 2 : (nb == 6 ? 3 : 4)));
    const int tid1 = fmt == W80211_F_L ? nbi : (two ? 9 + nbi : 4 + nbi);
    const uint16_t* map1 = T->map[tid1];
    const uint16_t* map2 = T->map[14 + nbi];
    float* out = llr + llr_off + (size_t)n * ncbps;
#pragma unroll
    for (int k2 = 0; k2 < 8; k2++) {
        int j = jj[k2];
        if (j >= 0) {
            cf32 q = c_divf(c_mul(e1[k2], ps), pa);
            qam_llr(q, nb, map1, j, out);
            if (two) {
                cf32 q2 = c_divf(c_mul(e2[k2], ps), pa);
                qam_llr(q2, nb, map2, j, out);
            }
        }
    }
}

void w_launch_demod(cudaStream_t st, const W80211Tables* tab, int nsymtot, int npkt, int nant, const float2* iq0,
                    const float2* iq1, const w80211_pkt_in* pk, const w80211_desc* desc, const int* sym_prefix,
                    const float2* chan, int chan_stride, float* llr)
{
    if (nsymtot <= 0) return;
    int per_block = DEMOD_THREADS / 8;
    int blocks = (nsymtot + per_block - 1) / per_block;
    demod_kernel<<<blocks, DEMOD_THREADS, 0, st>>>(tab, nsymtot, npkt, nant, iq0, iq1, pk, desc, sym_prefix, chan, chan_stride, llr);
}
