/* viterbi.cu -- soft-decision K=7 Viterbi decoder: forward pass and traceback, in two layouts.
 *
 * Replaces decode_impl::vstb_init / vstb_update / vstb_end (lib/decode_impl.cc:164-302).
 *
 * The reference's result is defined by one sequential FP32 add-compare-select chain per packet (no renormalisation,
 * strict '>' with the even source first, full-length traceback from state 0).  Both layouts keep exactly that chain per
 * state (two rounded candidate sums, their comparison, their maximum), so decisions and metrics are the reference's bit
 * for bit; overlapped windows would change which sums are formed and are not used.
 *
 *  - one THREAD per packet (viterbi_kernel, traceback_kernel): all 64 path metrics of a packet live in one thread's
 *    registers, every branch label is a compile-time constant (adds of an absent or zero branch metric disappear), no
 *    cross-lane traffic; the parallelism comes from the batch (32 packets per warp, 56 832 per wave).  Per state and step:
 *    <= 2 FADD (candidates), 1 FADD (difference, whose sign is the decision), 1 FMNMX, 1 SHF (decision bit into a word).
 *    Survivors: one 64-bit word per step and packet, bit d = 1 when destination state d took its odd source 2(d&31)+1,
 *    laid out [group of 32 packets][step pair][lane] as uint4 so a warp stores / loads 512 contiguous bytes.
 *  - one WARP per packet (viterbi_warp_kernel, traceback_warp_kernel, second half of this file): two metrics per lane,
 *    one shuffle exchange per step, survivors by ballot, exact parallel traceback.  For launches that cannot give every
 *    lane of the GPU a packet of its own (a single C4 packet: 1.3 ms instead of 7.4).
 * The host plan picks per code-rate group (capi.cu, w80211_cfg.flags).
 *
 * Depuncturing (decode_impl.cc:210-229) is folded into the step templates: a punctured position contributes exactly
 * 0.0f in the reference, i.e. the add is skipped here.
 */
#include "common.cuh"

#ifndef VIT_MINBLK
#define VIT_MINBLK 12 /* one-warp blocks per SM: 12 -> 168 registers, no spills; measured best (profiles/r1_viterbi_sweep.txt) */
#endif
/* trellis steps per unrolled loop body of the thread-per-packet kernel, per code rate: a whole number of puncturing
 * periods (1, 2, 3, 5 steps).  Measured on B200 (profiles/experiments/r2_viterbi_body_length.txt): six steps for rates 1/2, 2/3
 * and 3/4 (the state <-> register map closes by itself), five for rate 5/6 (closed by a register permutation) */
#ifndef VIT_STEPS_12
#define VIT_STEPS_12 6
#endif
#ifndef VIT_STEPS_23
#define VIT_STEPS_23 6
#endif
#ifndef VIT_STEPS_34
#define VIT_STEPS_34 6
#endif
#ifndef VIT_STEPS_56
#define VIT_STEPS_56 5
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

/* ---- in-place body of any whole number of puncturing periods.  After STEPS steps state x sits in register
 * rol6(x, STEPS mod 6); one permutation of the 64 registers closes the map (free in SASS: ptxas gives the last step's
 * results the registers the loop head expects).  This is what lets rate 5/6 (period 5, which does not divide 6) run in
 * place, and what lets every rate use a loop body that fits the instruction cache.  Numbers:
 * profiles/experiments/r2_viterbi_body_length.txt. */
template <int CR> __host__ __device__ constexpr int vit_period() { return CR == W80211_CR_12 ? 1 : (CR == W80211_CR_23 ? 2 : (CR == W80211_CR_34 ? 3 : 5)); }
__host__ __device__ constexpr int vit_mask_of(int k) { return k == 0 ? MASK_FULL : ((k & 1) ? MASK_A : MASK_B); } /* {1,1},{1,0},{0,1},{1,0},{0,1} */
__host__ __device__ constexpr int vit_llrs_before(int k) { return k == 0 ? 0 : k + 1; }                           /* inside one period */

template <int CR, int K>
__device__ __forceinline__ void ip_step(float (&m)[64], const float* v, uint2* base, int par)
{
    constexpr int PER = vit_period<CR>(), ph = K % PER, MASK = vit_mask_of(ph);
    constexpr int i0 = (K / PER) * (PER + 1) + vit_llrs_before(ph);
    const float t0 = MASK == MASK_B ? 0.f : v[i0];
    const float t1 = MASK == MASK_A ? 0.f : v[MASK == MASK_FULL ? i0 + 1 : i0];
    uint2 w;
    acs_ip<K % 6, MASK>(m, t0, t1, w.x, w.y);
    const int k = par + K; /* survivors: one uint4 per step pair and lane = two uint2, rows of 32 lanes */
    base[(size_t)(k >> 1) * 64 + (k & 1)] = w;
}

template <int CR, int K0, int N>
struct IpUnroll {
    static __device__ __forceinline__ void run(float (&m)[64], const float* v, uint2* base, int par)
    {
        ip_step<CR, K0>(m, v, base, par);
        IpUnroll<CR, K0 + 1, N - 1>::run(m, v, base, par);
    }
};
template <int CR, int K0>
struct IpUnroll<CR, K0, 0> {
    static __device__ __forceinline__ void run(float (&)[64], const float*, uint2*, int) {}
};

template <int NF>
__device__ __forceinline__ void ip_load(const float* __restrict__ llr, int b, float* v)
{
    /* body b starts at float b * NF of a 16-byte aligned packet */
    if (NF % 4 == 0) {
        const float4* src = reinterpret_cast<const float4*>(llr) + (size_t)b * (NF / 4);
#pragma unroll
        for (int k = 0; k < NF / 4; k++) {
            const float4 x = __ldg(src + k);
            v[4 * k] = x.x, v[4 * k + 1] = x.y, v[4 * k + 2] = x.z, v[4 * k + 3] = x.w;
        }
    } else if (NF % 2 == 0) {
        const float2* src = reinterpret_cast<const float2*>(llr) + (size_t)b * (NF / 2);
#pragma unroll
        for (int k = 0; k < NF / 2; k++) {
            const float2 x = __ldg(src + k);
            v[2 * k] = x.x, v[2 * k + 1] = x.y;
        }
    } else {
        const float* src = llr + (size_t)b * NF;
#pragma unroll
        for (int k = 0; k < NF; k++) v[k] = __ldg(src + k);
    }
}

template <int CR, int STEPS>
__device__ __forceinline__ void viterbi_run_ip(const float* __restrict__ llr, int trellis, uint4* __restrict__ surv4)
{
    constexpr int PER = vit_period<CR>(), NF = STEPS / PER * (PER + 1);
    static_assert(STEPS % PER == 0 && STEPS >= 1 && STEPS <= 12, "a body is a whole number of puncturing periods");
    const float NEG = -1000000000000000.0f; /* sentinel of vstb_init (decode_impl.cc:174-177) */
    float m[64];
#pragma unroll
    for (int i = 0; i < 64; i++) m[i] = NEG;
    m[0] = 0.0f;
    const int nbody = trellis / STEPS;
    uint2* surv2 = reinterpret_cast<uint2*>(surv4);
    float cur[NF], nxt[NF];
#pragma unroll
    for (int k = 0; k < NF; k++) cur[k] = 0.f;
    if (nbody > 0) ip_load<NF>(llr, 0, cur);
    constexpr int PFD = VIT_PF * 6 / STEPS; /* prefetch distance in bodies: about 48 steps ahead */
    for (int b = 0; b < nbody; b++) {
        ip_load<NF>(llr, b + 1 < nbody ? b + 1 : b, nxt);
        if (b + PFD < nbody) asm volatile("prefetch.global.L2 [%0];" ::"l"(llr + (size_t)(b + PFD) * NF));
        const int t0 = STEPS * b;
        IpUnroll<CR, 0, STEPS>::run(m, cur, surv2 + (size_t)(t0 >> 1) * 64, t0 & 1);
        if (STEPS % 6 != 0) {
            float n[64];
#pragma unroll
            for (int x = 0; x < 64; x++) n[x] = m[rol6(x, STEPS % 6)];
#pragma unroll
            for (int x = 0; x < 64; x++) m[x] = n[x];
        }
#pragma unroll
        for (int k = 0; k < NF; k++) cur[k] = nxt[k];
    }
    /* tail: fewer than STEPS steps; the map is the identity here, an absent LLR is an explicit 0.0f as in the reference */
    int t = nbody * STEPS, ph = 0;
    const float* q = llr + (size_t)nbody * NF;
    tail_ip<0>(m, t, trellis, ph, PER, q, surv2);
    if (STEPS > 2) tail_ip<1>(m, t, trellis, ph, PER, q, surv2);
    if (STEPS > 3) tail_ip<2>(m, t, trellis, ph, PER, q, surv2);
    if (STEPS > 4) tail_ip<3>(m, t, trellis, ph, PER, q, surv2);
    if (STEPS > 5) tail_ip<4>(m, t, trellis, ph, PER, q, surv2);
    if (STEPS > 6) tail_ip<5>(m, t, trellis, ph, PER, q, surv2);
    if (STEPS > 7) tail_ip<0>(m, t, trellis, ph, PER, q, surv2);
    if (STEPS > 8) tail_ip<1>(m, t, trellis, ph, PER, q, surv2);
    if (STEPS > 9) tail_ip<2>(m, t, trellis, ph, PER, q, surv2);
    if (STEPS > 10) tail_ip<3>(m, t, trellis, ph, PER, q, surv2);
    if (STEPS > 11) tail_ip<4>(m, t, trellis, ph, PER, q, surv2);
}

/* grp[g] = {x: first survivor row (even), y: rows reserved (even, >= max trellis in the group),
 *           z: first bit-word row, w: code rate of the group} */
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
    constexpr int STEPS = CR == W80211_CR_12 ? VIT_STEPS_12 : (CR == W80211_CR_23 ? VIT_STEPS_23 : (CR == W80211_CR_34 ? VIT_STEPS_34 : VIT_STEPS_56));
    if (STEPS == 6 && CR != W80211_CR_56)
        viterbi_run6<CR>(l, trellis, s4);
    else
        viterbi_run_ip<CR, STEPS>(l, trellis, s4);
}

/* vstb_end (decode_impl.cc:282-302): traceback from state 0 at the end of the trellis.  One thread
 * per packet; survivor words are read 16 steps (8 uint4, coalesced across the warp) at a time and
 * the decoded bits leave as one 32-bit word per 32 steps, bit i = decoded bit 32j+i.  64 registers, so
 * 32 one-warp blocks fit on an SM and a two-wave Viterbi batch is a single traceback wave. */
__global__ void __launch_bounds__(32, 32) traceback_kernel(const int4* __restrict__ grp, const int* __restrict__ slot_pkt,
                                                           const w80211_desc* __restrict__ desc, const uint2* __restrict__ surv,
                                                           uint32_t* __restrict__ bits, int g0)
{
    const int g = g0 + blockIdx.x, lane = threadIdx.x;
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

/* =====================================================================================================
 * Warp-per-packet variant, for launches that cannot fill the machine with one packet per lane.
 *
 * Same arithmetic per destination state as above (two FP32 candidate sums, strict '>' with the even
 * source first, max), so decisions and metrics are the reference's bit for bit; what changes is where
 * the 64 path metrics live: two per lane.  The state <-> (lane, register) map is not fixed: a
 * butterfly reads sources (2b, 2b+1) and writes (b, b+32) back into the two registers it read, and
 * before each step one __shfl_xor over lane bit j swaps one register between partner lanes so that
 * both sources of a butterfly sit in one lane.  j cycles 4,0,1,2,3 with the step number (t mod 5):
 *
 *     phase ph = t mod 5,  j = (ph + 4) mod 5,  lane holds butterfly b = rotr5(lane, ph)
 *     lanes with bit j clear hold even sources, their partners (bit j set) odd sources;
 *     send = bit j ? x0 : x1, after the exchange (ma, mb) = bit j ? (recv, x1) : (x0, recv);
 *     x0 <- max over sources for destination b, x1 <- for destination b + 32.
 *
 * State 0 is lane 0 / x0 in every phase, so the initial metrics need no special case.
 * Survivors: one uint2 per step and packet, (ballot of "odd source won" for x0, for x1) in LANE order
 * of that step; the traceback below walks (lane, register) positions, not state numbers, so the
 * permutation never has to be undone:
 *     decoded bit = r;  dec = bit `lane` of word r;  r' = bit j of lane;  lane' = lane with bit j := dec.
 * Layout: a group's survivor region (G.y rows of 32 uint2, as for the kernel above) is split into 32
 * contiguous runs of G.y uint2, one per packet.
 * ===================================================================================================== */
#define WV_FULL 0xffffffffu
__device__ __forceinline__ float bsel(uint32_t m, float a, float b) /* m ? a : b, m all ones or zero: one LOP3 */
{
    return __uint_as_float((__float_as_uint(a) & m) | (__float_as_uint(b) & ~m));
}

struct WvLane {
    uint32_t up[5]; /* all ones when this lane has bit j(ph) set */
    /* label L of the lane's butterfly (2b --0--> b) as 0.0f / 1.0f factors: fA = bit A of L (the branch uses t0), fB = bit B
     * (uses t1); gA, gB the same for the complementary label 3 - L */
    float fA[5], fB[5], gA[5], gB[5];
};

__device__ __forceinline__ void wv_setup(int lane, WvLane& W)
{
#pragma unroll
    for (int ph = 0; ph < 5; ph++) {
        const int j = (ph + 4) % 5;
        const int b = ((lane >> ph) | (lane << (5 - ph))) & 31; /* rotr5(lane, ph) */
        const int L = w_branch_label(2 * b, 0);
        W.up[ph] = (lane >> j) & 1 ? 0xffffffffu : 0u;
        W.fA[ph] = (L & 2) ? 1.0f : 0.0f;
        W.fB[ph] = (L & 1) ? 1.0f : 0.0f;
        W.gA[ph] = (L & 2) ? 0.0f : 1.0f;
        W.gB[ph] = (L & 1) ? 0.0f : 1.0f;
        /* opaque to the compiler, or it turns the selects below into ISETP + SEL / FSEL chains (it runs out of predicate
         * registers and re-derives them every step) */
        asm volatile("" : "+r"(W.up[ph]), "+f"(W.fA[ph]), "+f"(W.fB[ph]), "+f"(W.gA[ph]), "+f"(W.gB[ph]));
    }
}

/* Issue cost on B200 (profiles/r1_pipebench.txt): LOP3 / SEL / FSETP take two cycles of a scheduler's ALU pipe, FADD / FMUL /
 * FFMA / FMNMX one.  So the branch metrics are picked with exact multiplications by 0.0f / 1.0f on the FMA pipe
 * (t * 1 = t, t * 0 = +-0, and m + (+-0) = m for every metric m: the sums are the reference's bit for bit as long as the
 * LLRs are finite), and both registers are shuffled so that only the two receive-side selects remain. */
template <int PH, int MASK>
__device__ __forceinline__ uint2 wv_step(float& x0, float& x1, float t0, float t1, const WvLane& W)
{
    constexpr int J = (PH + 4) % 5;
    const uint32_t up = W.up[PH];
    const float r0 = __shfl_xor_sync(WV_FULL, x0, 1 << J);
    const float r1 = __shfl_xor_sync(WV_FULL, x1, 1 << J);
    const float ma = bsel(up, r1, x0); /* even source 2b */
    const float mb = bsel(up, x1, r0); /* odd source 2b+1 */
    /* branch metrics {0, t1, t0, t1+t0}[label] (decode_impl.cc:231-234): label L on 2b->b and 2b+1->b+32, 3-L on the other two */
    float bl, bc;
    if (MASK == MASK_FULL) {
        bl = __fmaf_rn(t0, W.fA[PH], __fmul_rn(t1, W.fB[PH]));
        bc = __fmaf_rn(t0, W.gA[PH], __fmul_rn(t1, W.gB[PH]));
    } else if (MASK == MASK_A) {
        bl = __fmul_rn(t0, W.fA[PH]);
        bc = __fmul_rn(t0, W.gA[PH]);
    } else {
        bl = __fmul_rn(t1, W.fB[PH]);
        bc = __fmul_rn(t1, W.gB[PH]);
    }
    const float c0 = __fadd_rn(ma, bl), c1 = __fadd_rn(mb, bc);
    const float e0 = __fadd_rn(ma, bc), e1 = __fadd_rn(mb, bl);
    x0 = fmaxf(c0, c1);
    x1 = fmaxf(e0, e1);
    uint2 w;
    w.x = __ballot_sync(WV_FULL, c1 > c0);
    w.y = __ballot_sync(WV_FULL, e1 > e0);
    return w;
}

/* steps per unrolled body (a multiple of 5 and of the puncturing period) and the LLRs it consumes */
template <int CR> struct WvBody;
template <> struct WvBody<W80211_CR_12> { enum { STEPS = 10, NF = 20, VW = 4 }; };
template <> struct WvBody<W80211_CR_23> { enum { STEPS = 20, NF = 30, VW = 2 }; };
template <> struct WvBody<W80211_CR_34> { enum { STEPS = 15, NF = 20, VW = 4 }; };
template <> struct WvBody<W80211_CR_56> { enum { STEPS = 10, NF = 12, VW = 4 }; };

/* puncturing (decode_impl.cc:197-229, SV_PUNC_* cloud80211phy.cc:1857-1860): which LLRs step k of a period has */
template <int CR> __host__ __device__ constexpr int wv_period() { return CR == W80211_CR_12 ? 1 : (CR == W80211_CR_23 ? 2 : (CR == W80211_CR_34 ? 3 : 5)); }
__host__ __device__ constexpr int wv_mask_of(int k) { return k == 0 ? MASK_FULL : ((k & 1) ? MASK_A : MASK_B); }
__host__ __device__ constexpr int wv_llrs_before(int k) { return k == 0 ? 0 : k + 1; } /* phase 0 takes two, every later phase one */

template <int CR, int K>
__device__ __forceinline__ void wv_body_step(float& x0, float& x1, const float* v, const WvLane& W, uint2* dst, bool store)
{
    constexpr int PER = wv_period<CR>();
    constexpr int ph = K % PER, MASK = wv_mask_of(ph);
    constexpr int base = (K / PER) * (PER + 1) + wv_llrs_before(ph);
    const float t0 = MASK == MASK_B ? 0.f : v[base];
    const float t1 = MASK == MASK_A ? 0.f : v[MASK == MASK_FULL ? base + 1 : base];
    const uint2 w = wv_step<K % 5, MASK>(x0, x1, t0, t1, W);
    if (store) dst[K] = w;
}

template <int CR, int K0, int N>
struct WvUnroll {
    static __device__ __forceinline__ void run(float& x0, float& x1, const float* v, const WvLane& W, uint2* dst, bool store)
    {
        wv_body_step<CR, K0>(x0, x1, v, W, dst, store);
        WvUnroll<CR, K0 + 1, N - 1>::run(x0, x1, v, W, dst, store);
    }
};
template <int CR, int K0>
struct WvUnroll<CR, K0, 0> {
    static __device__ __forceinline__ void run(float&, float&, const float*, const WvLane&, uint2*, bool) {}
};

template <int CR>
__device__ __forceinline__ void wv_load(const float* __restrict__ llr, int b, float* v)
{
    enum { NF = WvBody<CR>::NF, VW = WvBody<CR>::VW };
    if (VW == 4) {
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
        const float2* src = reinterpret_cast<const float2*>(llr) + (size_t)b * (NF / 2);
#pragma unroll
        for (int k = 0; k < NF / 2; k++) {
            const float2 x = __ldg(src + k);
            v[2 * k] = x.x;
            v[2 * k + 1] = x.y;
        }
    }
}

/* one tail step with run-time phase */
__device__ __forceinline__ uint2 wv_step_rt(int ph, float& x0, float& x1, float t0, float t1, const WvLane& W)
{
    switch (ph) {
    case 0: return wv_step<0, MASK_FULL>(x0, x1, t0, t1, W);
    case 1: return wv_step<1, MASK_FULL>(x0, x1, t0, t1, W);
    case 2: return wv_step<2, MASK_FULL>(x0, x1, t0, t1, W);
    case 3: return wv_step<3, MASK_FULL>(x0, x1, t0, t1, W);
    default: return wv_step<4, MASK_FULL>(x0, x1, t0, t1, W);
    }
}

#ifndef WV_MINBLK
#define WV_MINBLK 16
#endif
template <int CR>
__global__ void __launch_bounds__(32, WV_MINBLK) viterbi_warp_kernel(const int4* __restrict__ grp, const int* __restrict__ slot_pkt,
                                                                      const w80211_desc* __restrict__ desc, const float* __restrict__ llr,
                                                                      uint2* __restrict__ surv, int slot0)
{
    enum { STEPS = WvBody<CR>::STEPS, NF = WvBody<CR>::NF };
    const int slot = slot0 + blockIdx.x, lane = threadIdx.x;
    const int pkt = slot_pkt[slot];
    if (pkt < 0) return;
    const int4 G = grp[slot >> 5];
    const int trellis = desc[pkt].trellis;
    const float* __restrict__ l = llr + desc[pkt].llr_off;
    uint2* __restrict__ s = surv + (size_t)G.x * 32 + (size_t)(slot & 31) * G.y;
    WvLane W;
    wv_setup(lane, W);
    const float NEG = -1000000000000000.0f; /* vstb_init, decode_impl.cc:174-177 */
    float x0 = lane == 0 ? 0.0f : NEG, x1 = NEG;
    const bool store = lane == 0;
    const int nbody = trellis / STEPS;
    float cur[NF], nxt[NF];
    if (nbody > 0) {
        wv_load<CR>(l, 0, cur);
    } else {
#pragma unroll
        for (int k = 0; k < NF; k++) cur[k] = 0.f;
    }
    for (int b = 0; b < nbody; b++) {
        const int bn = b + 1 < nbody ? b + 1 : b;
        wv_load<CR>(l, bn, nxt);
        if (b + VIT_PF < nbody) asm volatile("prefetch.global.L2 [%0];" ::"l"(l + (size_t)(b + VIT_PF) * NF));
        WvUnroll<CR, 0, STEPS>::run(x0, x1, cur, W, s + (size_t)b * STEPS, store);
#pragma unroll
        for (int k = 0; k < NF; k++) cur[k] = nxt[k];
    }
    /* tail: fewer than STEPS steps; an absent LLR is an explicit 0.0f as in the reference */
    const int period = wv_period<CR>();
    const float* q = l + (size_t)nbody * NF;
    int ph = 0;
    for (int t = nbody * STEPS; t < trellis; t++) {
        const bool hasA = ph == 0 || (ph & 1), hasB = ph == 0 || !(ph & 1);
        float t0 = 0.0f, t1 = 0.0f;
        if (hasA) t0 = __ldg(q++);
        if (hasB) t1 = __ldg(q++);
        const uint2 w = wv_step_rt(t % 5, x0, x1, t0, t1, W);
        if (store) s[t] = w;
        if (++ph >= period) ph = 0;
    }
}

/* vstb_end (decode_impl.cc:282-302) for the layout above, one warp per packet, exact and parallel.
 *
 * The trellis is cut into 32 chunks of W steps (W a multiple of 32), lane c walks chunk c backwards.  Only the last
 * chunk knows its entry position (state 0 at the end of the trellis = lane 0, register 0); every other lane starts with
 * the SET of all 64 positions and propagates the set through the survivor words (a handful of bitwise operations per
 * step, see below).  Survivor paths merge: after a few dozen steps the set has one member, and from there on the walk
 * is the true path whatever the entry was.  So: (1) all lanes walk `m` steps in set mode, m a multiple of 32, until
 * every lane's set is a single position; (2) they walk the rest of their chunk as a chain and emit its decoded bits;
 * (3) each lane now knows the position at the bottom of its chunk, which is the entry of the chunk below: one shuffle
 * hands it down and the top m steps of every chunk are walked again as a chain from the true entry.  If some set
 * never shrinks to one member inside its chunk (decisions that map positions one-to-one for a thousand steps;
 * not seen with real LLRs), lane 0 walks the whole trellis alone: the result is exact either way.
 *
 * Set step, phase j: members whose decision bit is 0 came from the even source, 1 from the odd one; the predecessor of
 * (lane, r) is (lane with bit j := dec, r' = old bit j of lane).  With Z / O the members with dec 0 / 1 (both registers
 * folded) and Mj the lanes with bit j set:  S0' = (Z & ~Mj) | ((O & ~Mj) << 2^j),  S1' = ((Z & Mj) >> 2^j) | (O & Mj). */
struct TbPos {
    uint32_t l, r;
};
__device__ __forceinline__ void tb_chain_step(TbPos& p, uint2 w, int j, uint32_t& word, int bitpos)
{
    word |= p.r << bitpos;
    const uint32_t sel = p.r ? w.y : w.x;
    const uint32_t dec = (sel >> p.l) & 1u;
    const uint32_t m = 1u << j;
    p.r = (p.l >> j) & 1u;
    p.l = (p.l & ~m) | (dec << j);
}
__device__ __forceinline__ void tb_set_step(uint32_t& s0, uint32_t& s1, uint2 w, int j)
{
    const uint32_t Mj = j == 0 ? 0xAAAAAAAAu : (j == 1 ? 0xCCCCCCCCu : (j == 2 ? 0xF0F0F0F0u : (j == 3 ? 0xFF00FF00u : 0xFFFF0000u)));
    const uint32_t Z = (s0 & ~w.x) | (s1 & ~w.y), O = (s0 & w.x) | (s1 & w.y);
    const int sh = 1 << j;
    s0 = (Z & ~Mj) | ((O & ~Mj) << sh);
    s1 = ((Z & Mj) >> sh) | (O & Mj);
}

/* 16 steps t0+15 .. t0 (t0 a multiple of 16) of one lane's chunk; steps >= T do not exist */
template <bool SET>
__device__ __forceinline__ void tb_block16(const uint2* __restrict__ s, int t0, int T, TbPos& p, uint32_t& s0, uint32_t& s1, uint32_t& word)
{
    uint4 w[8];
    const uint4* s4 = reinterpret_cast<const uint4*>(s + t0);
#pragma unroll
    for (int k = 0; k < 8; k++)
        if (t0 + 2 * k < T) w[k] = __ldg(s4 + k);
    const int sh = t0 & 16;
    int ph = (t0 + 15) % 5;
#pragma unroll
    for (int k = 15; k >= 0; k--) {
        const int j = ph == 0 ? 4 : ph - 1;
        if (t0 + k < T) {
            const uint2 sw = (k & 1) ? make_uint2(w[k >> 1].z, w[k >> 1].w) : make_uint2(w[k >> 1].x, w[k >> 1].y);
            if (SET)
                tb_set_step(s0, s1, sw, j);
            else
                tb_chain_step(p, sw, j, word, sh + k);
        }
        ph = ph == 0 ? 4 : ph - 1;
    }
}

__global__ void __launch_bounds__(32, 16) traceback_warp_kernel(const int4* __restrict__ grp, const int* __restrict__ slot_pkt,
                                                                 const w80211_desc* __restrict__ desc, const uint2* __restrict__ surv,
                                                                 uint32_t* __restrict__ bits, int slot0)
{
    const int slot = slot0 + blockIdx.x, lane = threadIdx.x;
    const int pkt = slot_pkt[slot];
    if (pkt < 0) return;
    const int4 G = grp[slot >> 5];
    const int T = desc[pkt].trellis;
    const uint2* __restrict__ s = surv + (size_t)G.x * 32 + (size_t)(slot & 31) * G.y;
    uint32_t* __restrict__ out = bits + (size_t)G.z * 32 + (slot & 31);
    /* steps per chunk, a multiple of 32 with 32 * W >= T; at least 256 so that short packets use fewer, longer chunks
     * (survivor paths need a few dozen steps to merge, more at low SNR) */
    const int W = max(((T + 1023) >> 10) << 5, 256);
    const int lo = lane * W, hi = lo + W;  /* this lane's steps [lo, hi) */
    /* (1) set mode from the top of the chunk */
    uint32_t s0 = hi >= T ? 1u : 0xffffffffu, s1 = hi >= T ? 0u : 0xffffffffu;
    TbPos p = {0u, 0u};
    uint32_t word = 0;
    int m = 0;
    bool merged = false;
    while (m < W) {
        tb_block16<true>(s, hi - m - 16, T, p, s0, s1, word);
        tb_block16<true>(s, hi - m - 32, T, p, s0, s1, word);
        m += 32;
        merged = __popc(s0) + __popc(s1) == 1;
        if (__all_sync(WV_FULL, merged)) break;
    }
    const bool all_merged = __all_sync(WV_FULL, merged);
    if (all_merged) {
        /* (2) the rest of the chunk as a chain */
        p.r = s1 ? 1u : 0u;
        p.l = (uint32_t)__ffs((int)(s0 | s1)) - 1u;
        for (int t = hi - m - 32; t >= lo; t -= 32) {
            word = 0;
            tb_block16<false>(s, t + 16, T, p, s0, s1, word);
            tb_block16<false>(s, t, T, p, s0, s1, word);
            if (t < T) out[(size_t)(t >> 5) * 32] = word;
        }
        /* (3) the top m steps again, from the position the chunk above ended at */
        uint32_t el = __shfl_down_sync(WV_FULL, p.l, 1), er = __shfl_down_sync(WV_FULL, p.r, 1);
        if (lane == 31) el = er = 0u;
        p.l = el;
        p.r = er;
        for (int t = hi - 32; t >= hi - m; t -= 32) {
            word = 0;
            tb_block16<false>(s, t + 16, T, p, s0, s1, word);
            tb_block16<false>(s, t, T, p, s0, s1, word);
            if (t < T) out[(size_t)(t >> 5) * 32] = word;
        }
    } else if (lane == 0) {
        p.l = p.r = 0u;
        for (int t = ((T - 1) >> 5) << 5; t >= 0; t -= 32) {
            word = 0;
            tb_block16<false>(s, t + 16, T, p, s0, s1, word);
            tb_block16<false>(s, t, T, p, s0, s1, word);
            out[(size_t)(t >> 5) * 32] = word;
        }
    }
}

/* warp_mask: bit cr set = that code rate's groups use the warp-per-packet kernels (forward pass and traceback must agree:
 * the survivor layouts differ) */
int w_launch_viterbi(cudaStream_t st, const int* cr_first /* [5]: first group of each code rate, [4] = ngroups */, const int4* grp,
                     const int* slot_pkt, const w80211_desc* desc, const float* llr, uint2* surv, int warp_mask)
{
    int launches = 0;
    for (int cr = 0; cr < 4; cr++) {
        int n = cr_first[cr + 1] - cr_first[cr];
        if (n <= 0) continue;
        launches++;
        const int g0 = cr_first[cr];
        if ((warp_mask >> cr) & 1) {
            switch (cr) {
            case W80211_CR_12: viterbi_warp_kernel<W80211_CR_12><<<n * 32, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, g0 * 32); break;
            case W80211_CR_23: viterbi_warp_kernel<W80211_CR_23><<<n * 32, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, g0 * 32); break;
            case W80211_CR_34: viterbi_warp_kernel<W80211_CR_34><<<n * 32, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, g0 * 32); break;
            default: viterbi_warp_kernel<W80211_CR_56><<<n * 32, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, g0 * 32); break;
            }
            continue;
        }
        switch (cr) {
        case W80211_CR_12: viterbi_kernel<W80211_CR_12><<<n, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, g0); break;
        case W80211_CR_23: viterbi_kernel<W80211_CR_23><<<n, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, g0); break;
        case W80211_CR_34: viterbi_kernel<W80211_CR_34><<<n, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, g0); break;
        default: viterbi_kernel<W80211_CR_56><<<n, 32, 0, st>>>(grp, slot_pkt, desc, llr, surv, g0); break;
        }
    }
    return launches;
}

/* one launch per run of code rates that share a survivor layout */
int w_launch_traceback(cudaStream_t st, const int* cr_first, const int4* grp, const int* slot_pkt, const w80211_desc* desc,
                       const uint2* surv, uint32_t* bits, int warp_mask)
{
    int launches = 0;
    for (int cr = 0; cr < 4;) {
        const int w = (warp_mask >> cr) & 1;
        int c2 = cr + 1;
        while (c2 < 4 && ((warp_mask >> c2) & 1) == w) c2++;
        const int g0 = cr_first[cr], n = cr_first[c2] - g0;
        cr = c2;
        if (n <= 0) continue;
        launches++;
        if (w)
            traceback_warp_kernel<<<n * 32, 32, 0, st>>>(grp, slot_pkt, desc, surv, bits, g0 * 32);
        else
            traceback_kernel<<<n, 32, 0, st>>>(grp, slot_pkt, desc, surv, bits, g0);
    }
    return launches;
}

/* =====================================================================================================
 * w80211_viterbi_noalloc(): LLRs -> decoded trellis bits for callers that cannot allocate (CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * It calls no cudaMalloc / cudaFree / allocating routine: inputs and outputs are the caller's device
 * buffers, the packet table travels as a kernel argument, and the survivor words -- the scratch memory of the Viterbi
 * forward pass -- live in the file-scope array below, which every GPU that loads this library holds until exit
 * (W80211_NOALLOC_MAX_PKTS x 32784 x 8 B = 4.0 MiB).  The kernels are variants of viterbi_warp_kernel /
 * traceback_warp_kernel above (same device functions, same arithmetic, so the same bits as every other path); the
 * existing kernels and entry points are untouched.
 *
 * In host code `g_na_surv` is only a stand-in for the symbol, not a device address: the address is taken once per device,
 * with that device current, by cudaGetSymbolAddress() and handed to the kernels as an argument.
 * Every call shares the array.  Nothing in it has to start at zero: the forward kernel writes every survivor word the
 * traceback kernel uses (steps 0 .. trellis-1 of each packet) before it runs, on the same stream; words left over from an
 * earlier call lie beyond `trellis` and are never used.  Calls that overlap on the GPU (different streams, or several
 * host threads) race on the array: serialise them. */
#define NA_STRIDE 32784 /* survivor words reserved per packet: W80211_T_MAX rounded up to a step pair */
static __device__ __align__(256) uint2 g_na_surv[(size_t)W80211_NOALLOC_MAX_PKTS * NA_STRIDE];

struct NaParams {
    w80211_noalloc_pkt pk[W80211_NOALLOC_MAX_PKTS];
};

template <int CR>
__global__ void __launch_bounds__(32, WV_MINBLK) viterbi_warp_na_kernel(const NaParams P, const float* __restrict__ llr, uint2* __restrict__ surv)
{
    enum { STEPS = WvBody<CR>::STEPS, NF = WvBody<CR>::NF };
    const int p = blockIdx.x, lane = threadIdx.x;
    if (P.pk[p].cr != CR) return; /* one launch per code rate present, every block looks at its own packet */
    const int trellis = P.pk[p].trellis;
    const float* __restrict__ l = llr + P.pk[p].llr_off;
    uint2* __restrict__ s = surv + (size_t)p * NA_STRIDE;
    WvLane W;
    wv_setup(lane, W);
    const float NEG = -1000000000000000.0f;
    float x0 = lane == 0 ? 0.0f : NEG, x1 = NEG;
    const bool store = lane == 0;
    const int nbody = trellis / STEPS;
    float cur[NF], nxt[NF];
#pragma unroll
    for (int k = 0; k < NF; k++) cur[k] = 0.f;
    if (nbody > 0) wv_load<CR>(l, 0, cur);
    for (int b = 0; b < nbody; b++) {
        wv_load<CR>(l, b + 1 < nbody ? b + 1 : b, nxt);
        WvUnroll<CR, 0, STEPS>::run(x0, x1, cur, W, s + (size_t)b * STEPS, store);
#pragma unroll
        for (int k = 0; k < NF; k++) cur[k] = nxt[k];
    }
    const int period = wv_period<CR>();
    const float* q = l + (size_t)nbody * NF;
    int ph = 0;
    for (int t = nbody * STEPS; t < trellis; t++) {
        const bool hasA = ph == 0 || (ph & 1), hasB = ph == 0 || !(ph & 1);
        float t0 = 0.0f, t1 = 0.0f;
        if (hasA) t0 = __ldg(q++);
        if (hasB) t1 = __ldg(q++);
        const uint2 w = wv_step_rt(t % 5, x0, x1, t0, t1, W);
        if (store) s[t] = w;
        if (++ph >= period) ph = 0;
    }
}

/* traceback_warp_kernel with the packet table by value and packet-contiguous output words */
__global__ void __launch_bounds__(32, 16) traceback_warp_na_kernel(const NaParams P, const uint2* __restrict__ surv, uint32_t* __restrict__ bits)
{
    const int p = blockIdx.x, lane = threadIdx.x;
    const int T = P.pk[p].trellis;
    const uint2* __restrict__ s = surv + (size_t)p * NA_STRIDE;
    uint32_t* __restrict__ out = bits + P.pk[p].bits_off;
    const int W = max(((T + 1023) >> 10) << 5, 256);
    const int lo = lane * W, hi = lo + W;
    uint32_t s0 = hi >= T ? 1u : 0xffffffffu, s1 = hi >= T ? 0u : 0xffffffffu;
    TbPos pos = {0u, 0u};
    uint32_t word = 0;
    int m = 0;
    bool merged = false;
    while (m < W) {
        tb_block16<true>(s, hi - m - 16, T, pos, s0, s1, word);
        tb_block16<true>(s, hi - m - 32, T, pos, s0, s1, word);
        m += 32;
        merged = __popc(s0) + __popc(s1) == 1;
        if (__all_sync(WV_FULL, merged)) break;
    }
    if (__all_sync(WV_FULL, merged)) {
        pos.r = s1 ? 1u : 0u;
        pos.l = (uint32_t)__ffs((int)(s0 | s1)) - 1u;
        for (int t = hi - m - 32; t >= lo; t -= 32) {
            word = 0;
            tb_block16<false>(s, t + 16, T, pos, s0, s1, word);
            tb_block16<false>(s, t, T, pos, s0, s1, word);
            if (t < T) out[t >> 5] = word;
        }
        uint32_t el = __shfl_down_sync(WV_FULL, pos.l, 1), er = __shfl_down_sync(WV_FULL, pos.r, 1);
        if (lane == 31) el = er = 0u;
        pos.l = el;
        pos.r = er;
        for (int t = hi - 32; t >= hi - m; t -= 32) {
            word = 0;
            tb_block16<false>(s, t + 16, T, pos, s0, s1, word);
            tb_block16<false>(s, t, T, pos, s0, s1, word);
            if (t < T) out[t >> 5] = word;
        }
    } else if (lane == 0) {
        pos.l = pos.r = 0u;
        for (int t = ((T - 1) >> 5) << 5; t >= 0; t -= 32) {
            word = 0;
            tb_block16<false>(s, t + 16, T, pos, s0, s1, word);
            tb_block16<false>(s, t, T, pos, s0, s1, word);
            out[t >> 5] = word;
        }
    }
}

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>

namespace {
enum { NA_MAX_DEVICES = 64 };
std::mutex g_na_mutex;
uint2* g_na_ptr[NA_MAX_DEVICES]; /* device address of g_na_surv on each device, taken at that device's first call */

int na_fail(char* err, size_t errlen, int code, const char* fmt, ...)
{
    if (err && errlen) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, errlen, fmt, ap);
        va_end(ap);
    }
    return code;
}

/* LLRs a packet's trellis consumes (depuncturing of decode_impl.cc:210-229): phase 0 of a period takes two, later phases one */
long long na_llr_needed(int trellis, int cr)
{
    static const int per[4] = {1, 2, 3, 5};
    const long long full = trellis / per[cr], rem = trellis % per[cr];
    return full * (per[cr] + 1) + (rem > 0 ? rem + 1 : 0);
}
} // namespace

extern "C" int w80211_viterbi_noalloc(int32_t device, void* stream, int32_t npkt, const w80211_noalloc_pkt* pkts, const float* llr,
                                      uint64_t nllr, uint32_t* bits, uint64_t nwords, char* err, size_t errlen)
{
    if (err && errlen) err[0] = 0;
    if (npkt < 0 || (npkt > 0 && (!pkts || !llr || !bits)))
        return na_fail(err, errlen, W80211_E_ARG, "w80211_viterbi_noalloc: npkt %d with a missing pkts / llr / bits pointer", npkt);
    if (npkt > W80211_NOALLOC_MAX_PKTS)
        return na_fail(err, errlen, W80211_E_CAPACITY,
                       "w80211_viterbi_noalloc: %d packets in one call, the static scratch holds %d (W80211_NOALLOC_MAX_PKTS); split the call or use "
                       "w80211_decode_batch", npkt, W80211_NOALLOC_MAX_PKTS);
    if (device < 0 || device >= NA_MAX_DEVICES) return na_fail(err, errlen, W80211_E_ARG, "w80211_viterbi_noalloc: device %d out of range", device);
    if (((uintptr_t)llr & 15u) != 0) return na_fail(err, errlen, W80211_E_ARG, "w80211_viterbi_noalloc: llr must be 16-byte aligned");
    NaParams P;
    memset(&P, 0, sizeof(P));
    int present = 0;
    for (int p = 0; p < npkt; p++) {
        const w80211_noalloc_pkt& k = pkts[p];
        if (k.cr < 0 || k.cr > 3) return na_fail(err, errlen, W80211_E_ARG, "w80211_viterbi_noalloc: packet %d: code rate %d not in 0..3", p, k.cr);
        if (k.trellis < 1 || k.trellis > W80211_T_MAX)
            return na_fail(err, errlen, W80211_E_CAPACITY, "w80211_viterbi_noalloc: packet %d: trellis %d outside 1..%d (the scratch reserves %d survivor words per packet)",
                           p, k.trellis, W80211_T_MAX, NA_STRIDE);
        if ((k.llr_off & 3u) != 0) return na_fail(err, errlen, W80211_E_ARG, "w80211_viterbi_noalloc: packet %d: llr_off must be a multiple of 4 floats", p);
        if (k.llr_off > nllr || (uint64_t)na_llr_needed(k.trellis, k.cr) > nllr - k.llr_off)
            return na_fail(err, errlen, W80211_E_ARG, "w80211_viterbi_noalloc: packet %d: %lld LLRs from %llu do not fit in %llu", p,
                           na_llr_needed(k.trellis, k.cr), (unsigned long long)k.llr_off, (unsigned long long)nllr);
        if (k.bits_off > nwords || (uint64_t)((k.trellis + 31) >> 5) > nwords - k.bits_off)
            return na_fail(err, errlen, W80211_E_ARG, "w80211_viterbi_noalloc: packet %d: %d output words from %llu do not fit in %llu", p,
                           (k.trellis + 31) >> 5, (unsigned long long)k.bits_off, (unsigned long long)nwords);
        P.pk[p] = k;
        present |= 1 << k.cr;
    }
    if (npkt == 0) return W80211_OK;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return na_fail(err, errlen, W80211_E_CUDA, "w80211_viterbi_noalloc: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    uint2* surv = nullptr;
    {
        /* once per device, with that device current: the symbol's address on THIS device */
        std::lock_guard<std::mutex> lock(g_na_mutex);
        if (!g_na_ptr[device]) {
            void* ptr = nullptr;
            e = cudaGetSymbolAddress(&ptr, g_na_surv);
            if (e != cudaSuccess || !ptr) {
                cudaGetLastError();
                return na_fail(err, errlen, W80211_E_CUDA, "w80211_viterbi_noalloc: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(e));
            }
            g_na_ptr[device] = static_cast<uint2*>(ptr);
        }
        surv = g_na_ptr[device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (present & 1) viterbi_warp_na_kernel<W80211_CR_12><<<npkt, 32, 0, st>>>(P, llr, surv);
    if (present & 2) viterbi_warp_na_kernel<W80211_CR_23><<<npkt, 32, 0, st>>>(P, llr, surv);
    if (present & 4) viterbi_warp_na_kernel<W80211_CR_34><<<npkt, 32, 0, st>>>(P, llr, surv);
    if (present & 8) viterbi_warp_na_kernel<W80211_CR_56><<<npkt, 32, 0, st>>>(P, llr, surv);
    traceback_warp_na_kernel<<<npkt, 32, 0, st>>>(P, surv, bits);
    e = cudaGetLastError();
    if (e != cudaSuccess) return na_fail(err, errlen, W80211_E_CUDA, "w80211_viterbi_noalloc: kernel launch: %s", cudaGetErrorString(e));
    return W80211_OK;
}
