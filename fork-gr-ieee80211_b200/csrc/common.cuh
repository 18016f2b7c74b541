/* common.cuh -- shared device helpers of libw80211rx.
 *
 * Arithmetic policy.  Everything that feeds a decision (LLRs, path metrics) is
 * computed as an explicit sequence of IEEE-754 operations that mirrors what
 * the reference's C++ compiles to on x86-64 (no FMA contraction there):
 *   complex * complex  -> (ac - bd, ad + bc), four binary32 products, rounded one by one
 *                         (what g++ emits for std::complex<float>::operator*)
 *   complex / complex  -> libgcc __divsc3 (GCC >= 12): products, sums and the two
 *                         divisions in binary64, then rounded to binary32
 *   std::abs(complex)  -> glibc hypotf: (float)sqrt((double)x*x + (double)y*y)
 * so the CUDA path reproduces the oracle's floats bit for bit, not just within
 * the 1e-4 tolerance the spec allows.  The library is built with -fmad=false and
 * the helpers below use the _rn intrinsics anyway.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/w80211rx.h"
#include "fft64_canon.h"

#define W_WARP 32
#define W_NMAPS 19
#define W_MAPLEN 416
#ifndef W_DEMOD_SPC
#define W_DEMOD_SPC 36 /* most symbols per demod work item (one 8-lane group walks them in order); the host plan counts items with it */
#endif
#ifndef W_DEMOD_SPC2
#define W_DEMOD_SPC2 16 /* the same for two-stream packets (W80211_SPC2 in the environment overrides it: tuning) */
#endif
#define W_XPOW8 4224

/* Constant tables, generated on the host from the 802.11 formulas (tables.cc)
 * and kept in device global memory. */
struct W80211Tables {
    float pilot_p[128];        /* p_0..p_126 pilot polarity (802.11-2016 17.3.5.10); reference PILOT_P cloud80211phy.cc:160-168 */
    float ltf_nl[64];          /* HT/VHT-LTF sign per FFT bin, 0 on unused bins; reference LTF_NL_28_F_FLOAT :113-129 */
    int8_t sig_demap[64];      /* FFT bin -> coded-bit index of a 48-bit BPSK SIG symbol, -1 unused; reference FFT_NL_SIG_DEMAP :31 */
    int8_t bin_j_l[64];        /* FFT bin -> data subcarrier index (ascending frequency), legacy 48; -1 = null/pilot */
    int8_t bin_j_nl[64];       /* same, HT/VHT 52 */
    uint8_t sigb_deint[52];    /* reference mapDeintVhtSigB20 :1503 */
    uint8_t sigb_int[52];      /* reference mapIntelVhtSigB20 :1508 */
    /* interleaved LLR index -> index inside the symbol's nCBPS output block.
     * 0..3 legacy nbpscs 1,2,4,6; 4..8 HT/VHT one stream nbpscs 1,2,4,6,8;
     * 9..13 two streams, stream 1 (deinterleave + stream de-parse); 14..18 stream 2. */
    uint16_t map[W_NMAPS][W_MAPLEN];
    /* map[] again, indexed [FFT bin][bit], 16 bytes per bin: binmap[m][i][b] = map[m][j(i) * nb + b], or the symbol's LLR count
     * (the staging row's dump slot) for bins without data and bits the modulation does not have */
    uint16_t binmap_align_[4]; /* keeps binmap on a 16-byte boundary (checked in demod.cu) */
    uint16_t binmap[W_NMAPS][64][8];
    /* CRC-32 (reflected 0xEDB88320) for post.cu: slice-by-4 tables (crc_tab[0] is the byte table) and x^(2^k) mod P */
    uint32_t crc_tab[4][256];
    uint32_t crc_x2n[32];
    uint32_t crc_xpow8[W_XPOW8]; /* x^(8n) mod P, n = 0 .. W_XPOW8-1 bytes */
    uint32_t scr_seq[128][8];    /* descrambler: 160 output bits of x^7+x^4+1 started from 7-bit state s, bit k of word k/32; words 5..7 = 0 */
};

struct cf32 {
    float x, y;
};

__device__ __forceinline__ cf32 c_make(float x, float y)
{
    cf32 r;
    r.x = x;
    r.y = y;
    return r;
}
__device__ __forceinline__ cf32 c_add(cf32 a, cf32 b) { return c_make(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y)); }
__device__ __forceinline__ cf32 c_sub(cf32 a, cf32 b) { return c_make(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y)); }
__device__ __forceinline__ cf32 c_conj(cf32 a) { return c_make(a.x, -a.y); }
__device__ __forceinline__ cf32 c_neg(cf32 a) { return c_make(-a.x, -a.y); }
__device__ __forceinline__ cf32 c_scale(cf32 a, float f) { return c_make(__fmul_rn(a.x, f), __fmul_rn(a.y, f)); }
__device__ __forceinline__ cf32 c_divf(cf32 a, float f) { return c_make(__fdiv_rn(a.x, f), __fdiv_rn(a.y, f)); }
__device__ __forceinline__ cf32 c_mul(cf32 a, cf32 b)
{
    float ac = __fmul_rn(a.x, b.x), bd = __fmul_rn(a.y, b.y);
    float ad = __fmul_rn(a.x, b.y), bc = __fmul_rn(a.y, b.x);
    return c_make(__fsub_rn(ac, bd), __fadd_rn(ad, bc));
}
/* (a+ib)/(c+id) as libgcc's __divsc3 does it when binary64 hardware exists. */
__device__ __forceinline__ cf32 c_div(cf32 n, cf32 d)
{
    double a = n.x, b = n.y, c = d.x, dd = d.y;
    double denom = __dadd_rn(__dmul_rn(c, c), __dmul_rn(dd, dd));
    double x = __ddiv_rn(__dadd_rn(__dmul_rn(a, c), __dmul_rn(b, dd)), denom);
    double y = __ddiv_rn(__dsub_rn(__dmul_rn(b, c), __dmul_rn(a, dd)), denom);
    return c_make(__double2float_rn(x), __double2float_rn(y));
}
/* The same quotients with the divisor known in advance (one channel tap serves every symbol of a packet):
 * y = RN(1/b) once, then q0 = RN(a*y), e = a - b*q0 (exact in an FMA), q = RN(q0 + e*y).  With y the correctly
 * rounded reciprocal this is RN(a/b) for every a (Markstein; Brisebarre, Muller, Raina, "Accelerating correctly
 * rounded floating-point division when the divisor is known in advance", IEEE TC 2004) unless the significand of
 * b is all ones, which the callers test for and send to the plain division.  tools/divcheck.cu compares both
 * paths on random operands on the GPU. */
__device__ __forceinline__ bool d_recip_unsafe(double b) { return (__double2hiint(b) & 0x000fffff) == 0x000fffff && __double2loint(b) == -1; }
__device__ __forceinline__ bool f_recip_unsafe(float b) { return (__float_as_uint(b) & 0x007fffffu) == 0x007fffffu; }
__device__ __forceinline__ double d_div_known(double a, double b, double y)
{
    double q0 = __dmul_rn(a, y);
    double e = __fma_rn(-q0, b, a);
    return __fma_rn(e, y, q0);
}
__device__ __forceinline__ float f_div_known(float a, float b, float y)
{
    float q0 = __fmul_rn(a, y);
    float e = __fmaf_rn(-q0, b, a);
    return __fmaf_rn(e, y, q0);
}
/* c_div with denom = c*c + d*d and y = RN(1/denom) precomputed (same rounding as c_div: the products of two
 * binary32 values are exact in binary64, so fma(b, d, a*c) = RN(RN(a*c) + RN(b*d))) */
__device__ __forceinline__ cf32 c_div_known(cf32 n, cf32 d, double denom, double y)
{
    double a = n.x, b = n.y, c = d.x, dd = d.y;
    double nx = __fma_rn(b, dd, __dmul_rn(a, c));
    double ny = __fma_rn(-a, dd, __dmul_rn(b, c));
    return c_make(__double2float_rn(d_div_known(nx, denom, y)), __double2float_rn(d_div_known(ny, denom, y)));
}
/* std::abs(std::complex<float>) = hypotf */
__device__ __forceinline__ float c_abs(cf32 a)
{
    double x = a.x, y = a.y;
    return __double2float_rn(__dsqrt_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y))));
}

/* ---- 64-point FFT on 8 lanes (fft64_canon.h, n = 8*n1 + n2, k = k1 + 8*k2).
 * Lane g (0..7 inside its 8-lane group) loads x[8*n1+g], runs the column FFT,
 * twiddles, exchanges through shared memory and runs the row FFT.  On return
 * lane g holds X[g + 8*k2] in (re[k2], im[k2]).  xch: 8 rows of 9 float2 per
 * group (the pad keeps the 64-bit accesses conflict free). */
#define W_FFT_XCH_FLOAT2 72
/* per-lane twiddles W64^(g*k1), k1 = 1..7, read once per thread */
__device__ __forceinline__ void fft64_twiddles(int g, float* wr, float* wi)
{
#pragma unroll
    for (int k1 = 0; k1 < 8; k1++) {
        wr[k1] = C8_W64_RE[(g * k1) & 63];
        wi[k1] = C8_W64_IM[(g * k1) & 63];
    }
}
/* same, with lane g's inputs x[8*n1+g] already in (re[n1], im[n1]) */
__device__ __forceinline__ void fft64_lane8_regs(int g, float2* xch, float* re, float* im, const float* wr, const float* wi);
__device__ __forceinline__ void fft64_lane8(const float2* __restrict__ src, int g, float2* xch, float* re, float* im,
                                            const float* wr, const float* wi, bool valid)
{
#pragma unroll
    for (int n1 = 0; n1 < 8; n1++) {
        float2 v = valid ? __ldg(&src[8 * n1 + g]) : make_float2(0.f, 0.f);
        re[n1] = v.x;
        im[n1] = v.y;
    }
    fft64_lane8_regs(g, xch, re, im, wr, wi);
}
__device__ __forceinline__ void fft64_lane8_regs(int g, float2* xch, float* re, float* im, const float* wr, const float* wi)
{
    c8_fft8(re, im);
    if (g) { /* n2*k1 == 0 is left untouched, as in c8_fft64() */
#pragma unroll
        for (int k1 = 1; k1 < 8; k1++) {
            float ar = re[k1], ai = im[k1];
            re[k1] = C8_SUB(C8_MUL(ar, wr[k1]), C8_MUL(ai, wi[k1]));
            im[k1] = C8_ADD(C8_MUL(ar, wi[k1]), C8_MUL(ai, wr[k1]));
        }
    }
#pragma unroll
    for (int k1 = 0; k1 < 8; k1++) xch[g * 9 + k1] = make_float2(re[k1], im[k1]);
    __syncwarp();
#pragma unroll
    for (int n2 = 0; n2 < 8; n2++) {
        float2 v = xch[n2 * 9 + g];
        re[n2] = v.x;
        im[n2] = v.y;
    }
    __syncwarp();
    c8_fft8(re, im);
}

/* K=7 code of 802.11 (generators 133, 171 octal): 2-bit label A<<1|B of the branch that
 * leaves state `st` (newest bit in bit 5) with input bit b.  Equals the reference's
 * SV_STATE_OUTPUT[st][b] (cloud80211phy.cc:1877-1887); checked in tests/test_tables.py. */
__host__ __device__ constexpr int w_branch_label(int st, int b)
{
    int x1 = (st >> 5) & 1, x2 = (st >> 4) & 1, x3 = (st >> 3) & 1, x5 = (st >> 1) & 1, x6 = st & 1;
    int A = b ^ x2 ^ x3 ^ x5 ^ x6;
    int B = b ^ x1 ^ x2 ^ x3 ^ x6;
    return (A << 1) | B;
}

/* kernels (one translation unit each) */
void w_launch_header(cudaStream_t st, const W80211Tables* tab, int npkt, int nant, const float2* iq0, const float2* iq1,
                     unsigned long long nsamples, const w80211_pkt_in* pk, const float2* chan_in, w80211_desc* desc,
                     float2* chan_out, int chan_stride, int mu_pos);
int w_demod_init_device(); /* per-device kernel attributes, called by w80211_create() with the context's device current */
void w_launch_demod(cudaStream_t st, const W80211Tables* tab, int nsymtot, int npkt, int nant, const float2* iq0,
                    const float2* iq1, const w80211_pkt_in* pk, const w80211_desc* desc, const int* sym_prefix,
                    const float2* chan, int chan_stride, float* llr, int max_ncbps);
/* signal stage: per-flag L-SIG result and the copy list of accepted packets */
struct w_sig_acc {
    unsigned long long src; /* sync position in the raw stream */
    unsigned long long dst; /* first output sample */
    int nsamp;
    float rad;
    int sync; /* row of chan_all */
    int pad_;
};
/* detection stage: one trigger flag the sync block acts on, and what comes back */
struct w_det_event {
    long long rs;  /* run whose 0x02 flags are the last before the trigger: start (negative: it began before this buffer) */
    long long re;  /* ... end of the range to search (exclusive): min(run end, t) */
    long long pos; /* where the sync block's IDLE scan began */
    long long t;   /* the trigger flag's sample */
    int do_sync;   /* 0: only the conj / run maximum are wanted (state at the end of the buffer) */
    int pad_;
};
struct w_det_result {
    float2 conj;
    float run_max;
    int valid, mindex;
    float rad, snr, rssi;
};
int w_edges_blocks(unsigned long long n);
void w_launch_run_edges(cudaStream_t st, const float* ac, unsigned long long n, int prev_above0, unsigned int cap,
                        unsigned int* counters, unsigned int* starts, unsigned int* ends, void* blk);
void w_launch_sync_events(cudaStream_t st, int nev, const w_det_event* ev, const float* ac, const float2* conj, const float2* sig,
                          float conj_ac0, float2 conj0, w_det_result* res);
void w_launch_presiso(cudaStream_t st, const float2* x, unsigned long long n, float* ac, float2* cj);
void w_launch_lsig(cudaStream_t st, const W80211Tables* tab, int nsync, const w80211_sync* syncs, const float2* iq0,
                   unsigned long long nsamples, int4* res, float2* chan_all);
void w_launch_cfo_copy(cudaStream_t st, int npkt, int nant, int max_nsamp, const w_sig_acc* acc, const float2* in0,
                       const float2* in1, float2* out0, float2* out1, const float2* chan_all, float2* chan_out);
void w_launch_sc16_to_fc32(cudaStream_t st, int npkt, int nant, const w80211_pkt_in* pkts, const void* raw0, const void* raw1,
                           float2* out0, float2* out1, unsigned long long nsamples, int bytes_per_sample);
void w_launch_upload(cudaStream_t st, void* dst, const void* pinned_src, size_t bytes);
int w_launch_viterbi(cudaStream_t st, const int* cr_first, const int4* grp, const int* slot_pkt, const w80211_desc* desc,
                     const float* llr, uint2* surv, int warp_mask);
int w_launch_traceback(cudaStream_t st, const int* cr_first, const int4* grp, const int* slot_pkt, const w80211_desc* desc,
                       const uint2* surv, uint32_t* bits, int warp_mask);
void w_launch_post(cudaStream_t st, const W80211Tables* tab, int npkt_dec, const int4* grp, const int* pkt_slot, const int* dec_list,
                   const w80211_desc* desc, const uint32_t* bits, const uint32_t* psdu_off, uint8_t* out_bytes,
                   w80211_mpdu* mpdus, unsigned int mpdu_cap, unsigned int* counters);
