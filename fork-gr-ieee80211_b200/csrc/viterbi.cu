/* viterbi.cu -- soft-decision K=7 Viterbi decoder, one THREAD per packet.
 *
 * Replaces decode_impl::vstb_init / vstb_update / vstb_end (lib/decode_impl.cc:164-302).
 *
 * Why one thread per packet.  The reference's result is defined by one sequential FP32
 * add-compare-select chain per packet (no renormalisation, strict '>' with the even source
 * first, full-length traceback from state 0).  Splitting a packet into overlapped windows
 * changes the FP32 sums and is not provably bit-exact, and a warp-per-packet butterfly spends
 * more issue slots on shuffles, selects and ballots than on the ACS itself.  With the whole
 * trellis of a packet in one thread all 64 path metrics live in registers, every branch label
 * is a compile-time constant (adds of an absent or zero branch metric disappear), there is no
 * cross-lane traffic at all, and the parallelism comes from the batch: 32 packets per warp,
 * thousands of packets per launch.  Arithmetic per step and state: <= 2 FADD (candidates),
 * 1 FADD (difference, whose sign is the decision), 1 FMNMX, 1 SHF (decision bit into a word).
 *
 * Survivors: one 64-bit word per step and packet, bit d = 1 when destination state d took its
 * odd source 2(d&31)+1.  Layout [group of 32 packets][step pair][lane] as uint4 so that a warp
 * stores and the traceback kernel loads 512 contiguous bytes.
 *
 * Depuncturing (decode_impl.cc:210-229) is folded into the step template: a punctured
 * position contributes exactly 0.0f in the reference, i.e. the add is skipped here.
 */
#include "common.cuh"

#ifndef VIT_MINBLK
#define VIT_MINBLK 12 /* one-warp blocks per SM: 12 -> 168 registers, no spills; measured best (profiles/r1_viterbi_sweep.txt) */
#endif
#ifndef VIT_INPLACE
#define VIT_INPLACE 1
#endif
#ifndef VIT_PF
#define VIT_PF 8 /* L1 prefetch distance, in bodies */
#endif
#define MASK_FULL 0 /* both LLRs present */
#define MASK_A 1    /* only A (t0) present, t1 = 0 */
#define MASK_B 2    /* only B (t1) present, t0 = 0 */

template <int MASK>
__device__ __forceinline__ float add_label(float m, int l, float t0, float t1, float t3)
{
    /* branch metric table of the reference: {0, t1, t0, t1+t0}[label] (decode_impl.cc:231-234) */
    if (MASK == MASK_FULL) return l == 0 ? m : __fadd_rn(m, l == 1 ? t1 : (l == 2 ? t0 : t3));
    if (MASK == MASK_A) return (l & 2) ? __fadd_rn(m, t0) : m;
    return (l & 1) ? __fadd_rn(m, t1) : m;
}

template <int MASK>
__device__ __forceinline__ void acs(const float (&m)[64], float (&n)[64], float t0, float t1, uint32_t& lo, uint32_t& hi)
{
    const float t3 = MASK == MASK_FULL ? __fadd_rn(t1, t0) : 0.0f;
    lo = 0;
    hi = 0;
#pragma unroll
    for (int s = 31; s >= 0; s--) {
        const int L = w_branch_label(2 * s, 0); /* label of 2s --0--> s; the other three follow by complement */
        const float ma = m[2 * s], mb = m[2 * s + 1];
        /* destination s (input bit 0): even source first, as the reference visits sources in ascending order */
        float c0 = add_label<MASK>(ma, L, t0, t1, t3);
        float c1 = add_label<MASK>(mb, 3 - L, t0, t1, t3);
        n[s] = fmaxf(c0, c1);
        lo = __funnelshift_l(__float_as_uint(__fsub_rn(c0, c1)), lo, 1); /* sign(c0-c1) = (c1 > c0) */
        /* destination s+32 (input bit 1) */
        float e0 = add_label<MASK>(ma, 3 - L, t0, t1, t3);
        float e1 = add_label<MASK>(mb, L, t0, t1, t3);
        n[s + 32] = fmaxf(e0, e1);
        hi = __funnelshift_l(__float_as_uint(__fsub_rn(e0, e1)), hi, 1);
    }
}

/* two steps m -> n -> m and one 16-byte survivor store */
template <int M0, int M1>
__device__ __forceinline__ void acs2(float (&m)[64], float (&n)[64], float a0, float a1, float b0, float b1, uint4* dst)
{
    uint4 w;
    acs<M0>(m, n, a0, a1, w.x, w.y);
    acs<M1>(n, m, b0, b1, w.z, w.w);
    *dst = w;
}

template <int CR> struct Body;
template <> struct Body<W80211_CR_12> { enum { STEPS = 4, NV = 2 }; };
template <> struct Body<W80211_CR_23> { enum { STEPS = 8, NV = 3 }; };
template <> struct Body<W80211_CR_34> { enum { STEPS = 6, NV = 2 }; };
template <> struct Body<W80211_CR_56> { enum { STEPS = 10, NV = 3 }; };

/* one body = STEPS trellis steps fed by NV float4 loads; survivors for step pair k go to dst[k*32] */
template <int CR>
__device__ __forceinline__ void body(float (&m)[64], float (&n)[64], const float4* v, uint4* dst)
{
    if (CR == W80211_CR_12) {
        acs2<MASK_FULL, MASK_FULL>(m, n, v[0].x, v[0].y, v[0].z, v[0].w, dst);
        acs2<MASK_FULL, MASK_FULL>(m, n, v[1].x, v[1].y, v[1].z, v[1].w, dst + 32);
    } else if (CR == W80211_CR_23) { /* {1,1},{1,0} */
        acs2<MASK_FULL, MASK_A>(m, n, v[0].x, v[0].y, v[0].z, 0.f, dst);
        acs2<MASK_FULL, MASK_A>(m, n, v[0].w, v[1].x, v[1].y, 0.f, dst + 32);
        acs2<MASK_FULL, MASK_A>(m, n, v[1].z, v[1].w, v[2].x, 0.f, dst + 64);
        acs2<MASK_FULL, MASK_A>(m, n, v[2].y, v[2].z, v[2].w, 0.f, dst + 96);
    } else if (CR == W80211_CR_34) { /* {1,1},{1,0},{0,1} */
        acs2<MASK_FULL, MASK_A>(m, n, v[0].x, v[0].y, v[0].z, 0.f, dst);
        acs2<MASK_B, MASK_FULL>(m, n, 0.f, v[0].w, v[1].x, v[1].y, dst + 32);
        acs2<MASK_A, MASK_B>(m, n, v[1].z, 0.f, 0.f, v[1].w, dst + 64);
    } else { /* 5/6: {1,1},{1,0},{0,1},{1,0},{0,1} */
        acs2<MASK_FULL, MASK_A>(m, n, v[0].x, v[0].y, v[0].z, 0.f, dst);
        acs2<MASK_B, MASK_A>(m, n, 0.f, v[0].w, v[1].x, 0.f, dst + 32);
        acs2<MASK_B, MASK_FULL>(m, n, 0.f, v[1].y, v[1].z, v[1].w, dst + 64);
        acs2<MASK_A, MASK_B>(m, n, v[2].x, 0.f, 0.f, v[2].y, dst + 96);
        acs2<MASK_A, MASK_B>(m, n, v[2].z, 0.f, 0.f, v[2].w, dst + 128);
    }
}

/* ---- in-place variant.  A butterfly reads states (2b, 2b+1) and writes states (b, b+32); writing the
 * results back into the two registers just read turns the state <-> register map into a 6-bit
 * rotation per step, identity again after 6 steps.  Six steps are unrolled with the map as a
 * compile-time constant, so only 64 metric registers are live (no second array). */
__host__ __device__ constexpr int rol6(int x, int k) { return ((x << k) | (x >> (6 - k))) & 63; }

template <int PH, int MASK>
__device__ __forceinline__ void acs_ip(float (&m)[64], float t0, float t1, uint32_t& lo, uint32_t& hi)
{
    const float t3 = MASK == MASK_FULL ? __fadd_rn(t1, t0) : 0.0f;
    lo = 0;
    hi = 0;
#pragma unroll
    for (int b = 31; b >= 0; b--) {
        const int L = w_branch_label(2 * b, 0);
        const int p = rol6(2 * b, PH), q = rol6(2 * b + 1, PH);
        const float ma = m[p], mb = m[q];
        float c0 = add_label<MASK>(ma, L, t0, t1, t3);
        float c1 = add_label<MASK>(mb, 3 - L, t0, t1, t3);
        float e0 = add_label<MASK>(ma, 3 - L, t0, t1, t3);
        float e1 = add_label<MASK>(mb, L, t0, t1, t3);
        m[p] = fmaxf(c0, c1);
        lo = __funnelshift_l(__float_as_uint(__fsub_rn(c0, c1)), lo, 1);
        m[q] = fmaxf(e0, e1);
        hi = __funnelshift_l(__float_as_uint(__fsub_rn(e0, e1)), hi, 1);
    }
}

template <int PH, int M0, int M1>
__device__ __forceinline__ void acs2_ip(float (&m)[64], float a0, float a1, float b0, float b1, uint4* dst)
{
    uint4 w;
    acs_ip<PH, M0>(m, a0, a1, w.x, w.y);
    acs_ip<PH + 1, M1>(m, b0, b1, w.z, w.w);
    *dst = w;
}

/* six steps, rates 1/2 (12 LLRs), 3/4 (8 LLRs) and 2/3 (9 LLRs: three periods of {1,1},{1,0}) */
template <int CR>
__device__ __forceinline__ void body6(float (&m)[64], const float* v, uint4* dst)
{
    if (CR == W80211_CR_12) {
        acs2_ip<0, MASK_FULL, MASK_FULL>(m, v[0], v[1], v[2], v[3], dst);
        acs2_ip<2, MASK_FULL, MASK_FULL>(m, v[4], v[5], v[6], v[7], dst + 32);
        acs2_ip<4, MASK_FULL, MASK_FULL>(m, v[8], v[9], v[10], v[11], dst + 64);
    } else if (CR == W80211_CR_23) {
        acs2_ip<0, MASK_FULL, MASK_A>(m, v[0], v[1], v[2], 0.f, dst);
        acs2_ip<2, MASK_FULL, MASK_A>(m, v[3], v[4], v[5], 0.f, dst + 32);
        acs2_ip<4, MASK_FULL, MASK_A>(m, v[6], v[7], v[8], 0.f, dst + 64);
    } else {
        acs2_ip<0, MASK_FULL, MASK_A>(m, v[0], v[1], v[2], 0.f, dst);
        acs2_ip<2, MASK_B, MASK_FULL>(m, 0.f, v[3], v[4], v[5], dst + 32);
        acs2_ip<4, MASK_A, MASK_B>(m, v[6], 0.f, 0.f, v[7], dst + 64);
    }
}

template <int PH>
__device__ __forceinline__ void tail_ip(float (&m)[64], int& t, int trellis, int& ph, int period, const float*& q, uint2* surv2)
{
    if (t < trellis) {
        bool hasA = ph == 0 || (ph & 1), hasB = ph == 0 || !(ph & 1);
        float t0 = 0.0f, t1 = 0.0f;
        if (hasA) t0 = __ldg(q++);
        if (hasB) t1 = __ldg(q++);
        uint2 w;
        acs_ip<PH, MASK_FULL>(m, t0, t1, w.x, w.y);
        surv2[((size_t)(t >> 1) * 32) * 2 + (t & 1)] = w;
        if (++ph >= period) ph = 0;
        t++;
    }
}

/* floats of one six-step body; rates 1/2 and 3/4 read them as aligned float4, rate 2/3 (9 floats) as scalars */
template <int CR> struct Body6 { enum { NF = CR == W80211_CR_12 ? 12 : (CR == W80211_CR_23 ? 9 : 8) }; };

template <int CR>
__device__ __forceinline__ void load_body6(const float* __restrict__ llr, int b, float* v)
{
    enum { NF = Body6<CR>::NF };
    if (NF % 4 == 0) {
        const float4* src = reinterpret_cast<const float4*>(llr) + (size_t)b * (NF / 4);
#pragma unroll
        for (int k = 0; k < NF / 4; k++) {
            const float4 x = __ldg(src + k);
            v[4 * k] = x.x;
            v[4 * k + 1] = x.y;
            v[4 * k + 2] = x.z;
            v[4 * k + 3] = x.w;
        }
    } else {
        const float* src = llr + (size_t)b * NF;
#pragma unroll
        for (int k = 0; k < NF; k++) v[k] = __ldg(src + k);
    }
}

template <int CR>
__device__ __forceinline__ void viterbi_run6(const float* __restrict__ llr, int trellis, uint4* __restrict__ surv4)
{
    enum { NF = Body6<CR>::NF };
    const float NEG = -1000000000000000.0f;
    float m[64];
#pragma unroll
    for (int i = 0; i < 64; i++) m[i] = NEG;
    m[0] = 0.0f;
    const int nbody = trellis / 6;
    /* next body's LLRs are loaded into registers while the current body computes; L2 prefetch further ahead */
    float cur[NF], nxt[NF];
    if (nbody > 0) {
        load_body6<CR>(llr, 0, cur);
    } else {
#pragma unroll
        for (int k = 0; k < NF; k++) cur[k] = 0.f;
    }
    for (int b = 0; b < nbody; b++) {
        const int bn = b + 1 < nbody ? b + 1 : b;
        load_body6<CR>(llr, bn, nxt);
        if (b + VIT_PF < nbody) asm volatile("prefetch.global.L2 [%0];" ::"l"(llr + (size_t)(b + VIT_PF) * NF));
        body6<CR>(m, cur, surv4 + (size_t)b * 3 * 32);
#pragma unroll
        for (int k = 0; k < NF; k++) cur[k] = nxt[k];
    }
    int t = nbody * 6, ph = 0;
    const int period = CR == W80211_CR_12 ? 1 : (CR == W80211_CR_23 ? 2 : 3);
    const float* q = llr + (size_t)nbody * NF;
    uint2* surv2 = reinterpret_cast<uint2*>(surv4);
    tail_ip<0>(m, t, trellis, ph, period, q, surv2);
    tail_ip<1>(m, t, trellis, ph, period, q, surv2);
    tail_ip<2>(m, t, trellis, ph, period, q, surv2);
    tail_ip<3>(m, t, trellis, ph, period, q, surv2);
    tail_ip<4>(m, t, trellis, ph, period, q, surv2);
}

/* grp[g] = {x: first survivor row (even), y: rows reserved (even, >= max trellis in the group),
 *           z: first bit-word row, w: code rate of the group} */
template <int CR>
__device__ __forceinline__ void viterbi_run(const float* __restrict__ llr, int trellis, uint4* __restrict__ surv4 /* + lane */)
{
    enum { STEPS = Body<CR>::STEPS, NV = Body<CR>::NV };
    const float NEG = -1000000000000000.0f; /* sentinel of vstb_init (decode_impl.cc:174-177) */
    float m[64], n[64];
#pragma unroll
    for (int i = 0; i < 64; i++) m[i] = NEG;
    m[0] = 0.0f;

    const int nbody = trellis / STEPS;
    const float4* src = reinterpret_cast<const float4*>(llr);
    float4 cur[NV], nxt[NV];
    if (nbody > 0) {
#pragma unroll
        for (int k = 0; k < NV; k++) cur[k] = __ldg(src + k);
    }
    for (int b = 0; b < nbody; b++) {
        if (b + 1 < nbody) {
#pragma unroll
            for (int k = 0; k < NV; k++) nxt[k] = __ldg(src + (size_t)(b + 1) * NV + k);
        }
        body<CR>(m, n, cur, surv4 + (size_t)b * (STEPS / 2) * 32);
#pragma unroll
        for (int k = 0; k < NV; k++) cur[k] = nxt[k];
    }
    /* tail: fewer than STEPS steps left; same arithmetic through the full-step template with an
     * explicit 0.0f for a punctured position (x + 0.0f == x) */
    int t = nbody * STEPS;
    if (t < trellis) {
        const int period = CR == W80211_CR_12 ? 1 : (CR == W80211_CR_23 ? 2 : (CR == W80211_CR_34 ? 3 : 5));
        const float* q = llr + (size_t)nbody * NV * 4;
        uint2* surv2 = reinterpret_cast<uint2*>(surv4);
        int ph = 0;
        for (; t < trellis; t++) {
            /* puncture pattern pairs (A present, B present): phase 0 {1,1}; odd phases {1,0}; even > 0 {0,1} */
            bool hasA = ph == 0 || (ph & 1), hasB = ph == 0 || !(ph & 1);
            float t0 = 0.0f, t1 = 0.0f;
            if (hasA) t0 = __ldg(q++);
            if (hasB) t1 = __ldg(q++);
            uint2 w;
            acs<MASK_FULL>(m, n, t0, t1, w.x, w.y);
#pragma unroll
            for (int i = 0; i < 64; i++) m[i] = n[i];
            surv2[((size_t)(t >> 1) * 32) * 2 + (t & 1)] = w;
            if (++ph >= period) ph = 0;
        }
    }
}

/* one kernel per code rate (groups are sorted by rate), so each gets its own register allocation */
template <int CR>
__global__ void __launch_bounds__(32, VIT_MINBLK) viterbi_kernel(const int4* __restrict__ grp, const int* __restrict__ slot_pkt,
                                                                 const w80211_desc* __restrict__ desc, const float* __restrict__ llr,
                                                                 uint2* __restrict__ surv, int g0)
{
    const int g = g0 + blockIdx.x, lane = threadIdx.x;
    const int4 G = grp[g];
    const int pkt = slot_pkt[g * 32 + lane];
    if (pkt < 0) return;
    const int trellis = desc[pkt].trellis;
    const float* l = llr + desc[pkt].llr_off;
    uint4* s4 = reinterpret_cast<uint4*>(surv + (size_t)G.x * 32) + lane;
#if VIT_INPLACE
    if (CR == W80211_CR_12 || CR == W80211_CR_34 || CR == W80211_CR_23)
        viterbi_run6<CR>(l, trellis, s4);
    else
#endif
        viterbi_run<CR>(l, trellis, s4);
}

/* vstb_end (decode_impl.cc:282-302): traceback from state 0 at the end of the trellis.  One thread
 * per packet; survivor words are read 16 steps (8 uint4, coalesced across the warp) at a time and
 * the decoded bits leave as one 32-bit word per 32 steps, bit i = decoded bit 32j+i.  64 registers, so
 * 32 one-warp blocks fit on an SM and a two-wave Viterbi batch is a single traceback wave. */
__global__ void __launch_bounds__(32, 32) traceback_kernel(const int4* __restrict__ grp, const int* __restrict__ slot_pkt,
                                                           const w80211_desc* __restrict__ desc, const uint2* __restrict__ surv,
                                                           uint32_t* __restrict__ bits)
{
    const int g = blockIdx.x, lane = threadIdx.x;
    const int4 G = grp[g];
    const int pkt = slot_pkt[g * 32 + lane];
    if (pkt < 0) return;
    const int T = desc[pkt].trellis;
    const uint4* s4 = reinterpret_cast<const uint4*>(surv + (size_t)G.x * 32) + lane;
    uint32_t* out = bits + (size_t)G.z * 32 + lane;
    int st = 0;
    uint32_t word = 0;
    for (int c = (T - 1) >> 4; c >= 0; c--) { /* chunk c = steps 16c .. 16c+15 */
        uint4 w[8];
        const int t0 = c << 4;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (t0 + 2 * k < T) w[k] = __ldg(s4 + (size_t)((t0 >> 1) + k) * 32);
        }
        const int sh = (c & 1) << 4;
#pragma unroll
        for (int k = 7; k >= 0; k--) {
            /* step t0+2k+1 (z,w) then t0+2k (x,y) */
            if (t0 + 2 * k + 1 < T) {
                uint32_t sel = (st & 32) ? w[k].w : w[k].z;
                word |= (uint32_t)(st >> 5) << (sh + 2 * k + 1);
                st = ((st & 31) << 1) | ((sel >> (st & 31)) & 1u);
            }
            if (t0 + 2 * k < T) {
                uint32_t sel = (st & 32) ? w[k].y : w[k].x;
                word |= (uint32_t)(st >> 5) << (sh + 2 * k);
                st = ((st & 31) << 1) | ((sel >> (st & 31)) & 1u);
            }
        }
        if (!(c & 1)) {
            out[(size_t)(c >> 1) * 32] = word;
            word = 0;
        }
    }
}

int w_launch_viterbi(cudaStream_t st, const int* cr_first /* [5]: first group of each code rate, [4] = ngroups */, const int4* grp,
                     const int* slot_pkt, const w80211_desc* desc, const float* llr, uint2* surv)
{
    int launches = 0;
    for (int cr = 0; cr < 4; cr++) {
        int n = cr_first[cr + 1] - cr_first[cr];
        if (n <= 0) continue;
        launches++;
        switch (cr) {
        case W80211_CR_12: viterbi_kernel<W80211_CR_12><<<n, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, cr_first[cr]); break;
        case W80211_CR_23: viterbi_kernel<W80211_CR_23><<<n, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, cr_first[cr]); break;
        case W80211_CR_34: viterbi_kernel<W80211_CR_34><<<n, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, cr_first[cr]); break;
        default: viterbi_kernel<W80211_CR_56><<<n, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, cr_first[cr]); break;
        }
    }
    return launches;
}

void w_launch_traceback(cudaStream_t st, int ngroups, const int4* grp, const int* slot_pkt, const w80211_desc* desc,
                        const uint2* surv, uint32_t* bits)
{
    if (ngroups <= 0) return;
    traceback_kernel<<<ngroups, 32, 0, st>>>(grp, slot_pkt, desc, surv, bits);
}
