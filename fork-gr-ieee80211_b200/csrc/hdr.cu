/* hdr.cu -- per-packet header stage, one warp per packet.
 *
 * Replaces the RDTAG / FORMAT / VHT / HT / LEGACY states of the reference's demod and demod2
 * blocks (lib/demod_impl.cc:72-219, lib/demod2_impl.cc:70-225): format detection from the two
 * symbols after L-SIG (procNLSigDemodDeint cloud80211phy.cc:629-648, svSigDecoder::decode
 * :2001-2088, signalCheckVhtA :753-771, signalCheckHt :730-751), SIG parsing (signalParserL
 * :773-850, signalParserHt :852-998, signalParserVhtA :1090-1178, signalParserVhtB :1180-1223,
 * modParserVht :1225-1323), the HT/VHT-LTF channel estimate (demod_impl.cc:380-392,
 * demod2_impl.cc:356-464), VHT-SIG-B (demod_impl.cc:449-505, demod2_impl.cc:632-758) and the
 * length / duration gates (demod_impl.cc:159,186; demod2_impl.cc:159,186).
 * Output: one w80211_desc per packet and the per-packet channel block the demod kernel reads.
 */
#include "common.cuh"

struct HdrSmem {
    float2 xch[4][W_FFT_XCH_FLOAT2];
    float2 F[4][64];     /* FFT outputs, natural bin order */
    float2 eq[2][64];    /* equalised bins (d_sig1 / d_sig2 of the reference) */
    float2 qam[2][52];   /* VHT-SIG-B symbols after phase correction */
    float llr[2][96];    /* [0] HT hypothesis, [1] VHT hypothesis; reused for SIG-B */
    uint8_t surv[49][64];
    uint8_t bits[48];
    uint8_t coded[52];
};

/* svSigDecoder::decode (cloud80211phy.cc:2001-2088): K=7 soft Viterbi over n <= 48 steps with the
 * reference's sentinel / strict '>' / ascending-source-order rules, survivors kept as source state
 * indices, full traceback from state 0.  Lane l owns destinations l and l+32, whose sources are
 * 2l and 2l+1 in that order. */
__device__ void sig_viterbi(HdrSmem& S, const float* llr, int n, int lane)
{
    const float NEG = -1000000000000000.0f;
    /* Path metrics stay in registers: lane l owns states l (lo) and l + 32 (hi) and fetches its two sources 2l, 2l + 1
     * with shuffles (lanes (2l) & 31 and (2l + 1) & 31; their lo half when l < 16, else their hi half) -- no shared-memory
     * round trip and no warp barrier per step. */
    float lo = lane == 0 ? 0.0f : NEG, hi = NEG;
    S.surv[0][lane] = 0;
    S.surv[0][lane + 32] = 0;
    const int L = w_branch_label(2 * lane, 0);
    /* the lane's two branch metrics {0, t1, t0, t1+t0}[L] and [3 - L], picked with masks made once: a select on the
     * lane-dependent label inside the loop compiles to branches, a quarter of this kernel's stall samples */
    const uint32_t a1 = L == 1 ? ~0u : 0u, a0 = L == 2 ? ~0u : 0u, a3 = L == 3 ? ~0u : 0u;
    const uint32_t b1 = L == 2 ? ~0u : 0u, b0 = L == 1 ? ~0u : 0u, b3 = L == 0 ? ~0u : 0u;
    const int sa = (2 * lane) & 31, sb = (2 * lane + 1) & 31;
    const bool low = lane < 16;
    __syncwarp(); /* the LLRs were written by other lanes */
    for (int t = 0; t < n; t++) {
        float t0 = llr[2 * t], t1 = llr[2 * t + 1];
        float t3 = __fadd_rn(t1, t0);
        const float alo = __shfl_sync(0xffffffffu, lo, sa), ahi = __shfl_sync(0xffffffffu, hi, sa);
        const float blo = __shfl_sync(0xffffffffu, lo, sb), bhi = __shfl_sync(0xffffffffu, hi, sb);
        float ma = low ? alo : ahi, mb = low ? blo : bhi;
        const uint32_t u0 = __float_as_uint(t0), u1 = __float_as_uint(t1), u3 = __float_as_uint(t3);
        float la = __uint_as_float((u1 & a1) | (u0 & a0) | (u3 & a3)), lb = __uint_as_float((u1 & b1) | (u0 & b0) | (u3 & b3));
        /* destination `lane` (input bit 0) */
        float c0 = __fadd_rn(ma, la), c1 = __fadd_rn(mb, lb);
        float best = NEG;
        int sv = 0;
        if (c0 > best) { best = c0; sv = 2 * lane; }
        if (c1 > best) { best = c1; sv = 2 * lane + 1; }
        lo = best;
        S.surv[t + 1][lane] = (uint8_t)sv;
        /* destination lane+32 (input bit 1): labels swap */
        c0 = __fadd_rn(ma, lb);
        c1 = __fadd_rn(mb, la);
        best = NEG;
        sv = 0;
        if (c0 > best) { best = c0; sv = 2 * lane; }
        if (c1 > best) { best = c1; sv = 2 * lane + 1; }
        hi = best;
        S.surv[t + 1][lane + 32] = (uint8_t)sv;
    }
    __syncwarp();
    if (lane == 0) {
        int st = 0;
        for (int j = n; j > 0; j--) {
            int prev = S.surv[j][st];
            /* bit j-1 = 1 iff st is the '1' successor of prev */
            S.bits[j - 1] = (st == ((prev >> 1) | 32)) ? 1 : 0;
            st = prev;
        }
    }
    __syncwarp();
}

/* checkBitCrc8 (cloud80211phy.cc:1367-1403): x^8+x^2+x+1, register preset to ones, inverted on air */
__device__ bool crc8_ok(const uint8_t* b)
{
    unsigned c = 0xff;
    for (int i = 0; i < 34; i++) {
        unsigned fb = ((c >> 7) & 1u) ^ b[i];
        c = ((c << 1) & 0xffu) ^ (fb ? 7u : 0u);
    }
    for (int i = 0; i < 8; i++)
        if (b[34 + i]) c ^= 1u << (7 - i);
    return (c & 0xffu) == 0xffu;
}

__device__ int bits_le(const uint8_t* b, int first, int n)
{
    int v = 0;
    for (int i = 0; i < n; i++) v |= ((int)b[first + i]) << i;
    return v;
}

/* modulation / coding of HT (mcs mod 8) and VHT MCS 0..9 */
__device__ void set_mod_nl(w80211_desc& d, int m)
{
    const int nb[10] = {1, 2, 2, 4, 4, 6, 6, 6, 8, 8};
    const int cr[10] = {W80211_CR_12, W80211_CR_12, W80211_CR_34, W80211_CR_12, W80211_CR_34,
                        W80211_CR_23, W80211_CR_34, W80211_CR_56, W80211_CR_34, W80211_CR_56};
    const int md[10] = {0, 2, 2, 3, 3, 4, 4, 4, 5, 5}; /* C8P_QAM_* (cloud80211phy.h:51-56) */
    d.nbpscs = nb[m];
    d.cr = cr[m];
    d.mod = md[m];
    d.ncbpss = d.nbpscs * 52;
    d.ncbps = d.ncbpss * d.nss;
    int num = d.cr == W80211_CR_12 ? 1 : (d.cr == W80211_CR_23 ? 2 : (d.cr == W80211_CR_34 ? 3 : 5));
    int den = d.cr == W80211_CR_12 ? 2 : (d.cr == W80211_CR_23 ? 3 : (d.cr == W80211_CR_34 ? 4 : 6));
    d.ndbps = (d.ncbps * num) / den;
    d.nltf = d.nss == 1 ? 1 : (d.nss == 2 ? 2 : 4);
}

__device__ void set_mod_legacy(w80211_desc& d, int rate, int len)
{
    const int nb[8] = {1, 1, 2, 2, 4, 4, 6, 6};
    const int cr[8] = {W80211_CR_12, W80211_CR_34, W80211_CR_12, W80211_CR_34, W80211_CR_12, W80211_CR_34, W80211_CR_23, W80211_CR_34};
    const int md[8] = {0, 0, 2, 2, 3, 3, 4, 4};
    const int nd[8] = {24, 36, 48, 72, 96, 144, 192, 216};
    d.format = W80211_F_L;
    d.mcs = rate;
    d.len = len;
    d.nbpscs = nb[rate];
    d.cr = cr[rate];
    d.mod = md[rate];
    d.ndbps = nd[rate];
    d.ncbps = d.ncbpss = 48 * d.nbpscs;
    d.nss = 1;
    d.nltf = 0;
    d.nsymsamp = 80;
    d.ampdu = 0;
    d.nsym = (len * 8 + 22 + d.ndbps - 1) / d.ndbps;
    d.trellis = len * 8 + 22;
    d.data_off = 0;
}

__global__ void __launch_bounds__(32) hdr_kernel(const W80211Tables* __restrict__ T, int npkt, int nant,
                                                 const float2* __restrict__ iq0, const float2* __restrict__ iq1,
                                                 unsigned long long nsamples, const w80211_pkt_in* __restrict__ pkin,
                                                 const float2* __restrict__ chan_in, w80211_desc* __restrict__ desc_out,
                                                 float2* __restrict__ chan_out, int chan_stride, int mu_pos)
{
    __shared__ HdrSmem S;
    const int p = blockIdx.x;
    if (p >= npkt) return;
    const int lane = threadIdx.x;
    const int grp = lane >> 3, g = lane & 7;
    const w80211_pkt_in pk = pkin[p];
    const float2* hl = chan_in + (size_t)p * 64;
    float2* cho = chan_out + (size_t)p * chan_stride;

    w80211_desc d;
    memset(&d, 0, sizeof(d));
    d.cfo = pk.cfo;
    d.snr = pk.snr;
    d.rssi = pk.rssi;
    d.l_mcs = pk.l_mcs;
    d.llr_off = -1;

    bool ok_in = pk.l_mcs >= 0 && pk.l_mcs <= 7 && pk.l_len >= 0 && pk.l_len <= 4095 && pk.nsamp >= 0 &&
                 pk.iq_offset + (unsigned long long)pk.nsamp <= nsamples;
    if (!ok_in) {
        d.status = W80211_ST_BAD_INPUT;
        if (lane == 0) desc_out[p] = d;
        return;
    }
    const float2* x0 = iq0 + pk.iq_offset;
    const float2* x1 = nant == 2 ? iq1 + pk.iq_offset : x0;

    int fmt = W80211_F_L;
    if (pk.l_mcs == 0 && pk.nsamp >= 160) {
        /* ---------------- FORMAT (demod_impl.cc:106-148) ---------------- */
        float wr[8], wi[8], re[8], im[8];
        fft64_twiddles(g, wr, wi);
        /* group 0: symbol at +8, group 1: symbol at +88, group 2/3: first HT/VHT-LTF of antenna 0/1 at +248 */
        const float2* src = grp == 0 ? x0 + 8 : (grp == 1 ? x0 + 88 : (grp == 2 ? x0 + 248 : x1 + 248));
        bool valid = grp < 2 || pk.nsamp >= 320;
        fft64_lane8(src, g, S.xch[grp], re, im, wr, wi, valid);
#pragma unroll
        for (int k2 = 0; k2 < 8; k2++) S.F[grp][g + 8 * k2] = make_float2(re[k2], im[k2]);
        __syncwarp();

        /* procNLSigDemodDeint: common phase of each symbol from the four pilots (+,-,+,+) */
        cf32 ps[2];
        float pa[2];
#pragma unroll
        for (int s = 0; s < 2; s++) {
            cf32 e7 = c_div(c_make(S.F[s][7].x, S.F[s][7].y), c_make(hl[7].x, hl[7].y));
            cf32 e21 = c_div(c_make(S.F[s][21].x, S.F[s][21].y), c_make(hl[21].x, hl[21].y));
            cf32 e43 = c_div(c_make(S.F[s][43].x, S.F[s][43].y), c_make(hl[43].x, hl[43].y));
            cf32 e57 = c_div(c_make(S.F[s][57].x, S.F[s][57].y), c_make(hl[57].x, hl[57].y));
            ps[s] = c_conj(c_add(c_add(c_sub(e7, e21), e43), e57));
            pa[s] = c_abs(ps[s]);
        }
#pragma unroll
        for (int h = 0; h < 2; h++) {
            int i = lane + 32 * h;
            int dm = T->sig_demap[i];
            if (dm >= 0) {
                cf32 hh = c_make(hl[i].x, hl[i].y);
                cf32 m1 = c_divf(c_mul(c_div(c_make(S.F[0][i].x, S.F[0][i].y), hh), ps[0]), pa[0]);
                cf32 m2 = c_divf(c_mul(c_div(c_make(S.F[1][i].x, S.F[1][i].y), hh), ps[1]), pa[1]);
                S.llr[0][dm] = m1.y;      /* HT-SIG: QBPSK on both symbols */
                S.llr[0][dm + 48] = m2.y;
                S.llr[1][dm] = m1.x;      /* VHT-SIG-A: BPSK then QBPSK */
                S.llr[1][dm + 48] = m2.y;
            }
        }
        __syncwarp();
        sig_viterbi(S, S.llr[1], 48, lane);
        const uint8_t* b = S.bits;
        bool vht = (b[2] == 1) && (b[23] == 1) && (b[33] == 1) && crc8_ok(b) && (b[0] + b[1] == 0);
        if (vht) {
            fmt = W80211_F_VHT;
        } else {
            __syncwarp();
            sig_viterbi(S, S.llr[0], 48, lane);
            bool ht = (b[26] == 1) && crc8_ok(b) && (b[5] + b[6] + b[7] + b[28] + b[29] + b[30] + b[32] + b[33] == 0);
            if (ht) fmt = W80211_F_HT;
        }
    }

    if (fmt == W80211_F_L) {
        /* LEGACY state (demod_impl.cc:207-219) */
        if (pk.l_len <= 0) {
            d.status = W80211_ST_BAD_INPUT;
        } else {
            set_mod_legacy(d, pk.l_mcs, pk.l_len);
            if ((long long)d.nsym * 80 > (long long)pk.nsamp) d.status = W80211_ST_BAD_INPUT;
        }
        for (int i = lane; i < 64; i += 32) cho[i] = hl[i];
        d.total = d.nsym * d.ncbps;
        if (lane == 0) desc_out[p] = d;
        return;
    }

    const uint8_t* b = S.bits;
    int sgi;
    bool mu = false;       /* VHT MU PPDU on the one-antenna block (demod_impl.cc:347-379) */
    bool sounding = false; /* two-stream VHT packet on the one-antenna block: only an NDP is meaningful (demod_impl.cc:393-409) */
    d.format = fmt;
    if (fmt == W80211_F_HT) {
        /* signalParserHt */
        int mcs = bits_le(b, 0, 7);
        d.mcs = mcs;
        d.len = bits_le(b, 8, 16);
        d.ampdu = b[27];
        sgi = b[31];
        d.nss = mcs / 8 + 1;
        set_mod_nl(d, mcs % 8);
        d.nsym = (d.len * 8 + 22 + d.ndbps - 1) / d.ndbps;
        d.trellis = d.len * 8 + 22;
    } else {
        /* signalParserVhtA */
        int gid = bits_le(b, 4, 6);
        sgi = b[24];
        d.ampdu = 1;
        if (gid != 0 && gid != 63) {
            /* MU: one stream for this user, two VHT-LTFs; modulation and length come from VHT-SIG-B
             * (signalParserVhtA cloud80211phy.cc:1166-1173).  demod2 has no MU branch. */
            if (nant == 2) {
                d.status = W80211_ST_UNSUPPORTED;
            } else {
                mu = true;
                d.nss = 1;
                set_mod_nl(d, 0);
                d.nltf = 2;
            }
        } else {
            d.nss = bits_le(b, 10, 3) + 1;
            d.mcs = bits_le(b, 28, 4);
            if (d.nss > 2)
                d.status = W80211_ST_DROP_GATE;
            else if (d.mcs > 9)
                d.status = W80211_ST_UNSUPPORTED; /* reference keeps stale modulation fields here */
            else
                set_mod_nl(d, d.mcs);
        }
    }
    d.nsymsamp = sgi ? 72 : 80;
    if (fmt == W80211_F_HT && d.nss > 2) d.status = W80211_ST_DROP_GATE;
    if (d.status == W80211_ST_OK && nant == 1 && d.nss == 2) {
        if (fmt == W80211_F_VHT)
            sounding = true;
        else
            d.status = W80211_ST_UNSUPPORTED;
    }
    const int hdr_len = 160 + 80 + 80 * d.nltf + (fmt == W80211_F_VHT ? 80 : 0);
    if (d.status == W80211_ST_OK && pk.nsamp < hdr_len) d.status = W80211_ST_DROP_GATE;
    if (d.status != W80211_ST_OK) {
        if (lane == 0) desc_out[p] = d;
        return;
    }

    /* ---------------- non-legacy channel estimate ---------------- */
    const int sigb_at = 240 + 80 * d.nltf + 8;
    if (d.nss == 1 || sounding) {
        if (fmt == W80211_F_VHT) {
            /* group 0: VHT-SIG-B; group 1: the second VHT-LTF (MU only) */
            float wr[8], wi[8], re[8], im[8];
            fft64_twiddles(g, wr, wi);
            fft64_lane8(grp == 0 ? x0 + sigb_at : x0 + 328, g, S.xch[grp], re, im, wr, wi, grp == 0 || (grp == 1 && mu));
            if (grp < 2) {
#pragma unroll
                for (int k2 = 0; k2 < 8; k2++) S.F[grp][g + 8 * k2] = make_float2(re[k2], im[k2]);
            }
            __syncwarp();
        }
        /* H = FFT(LTF) / L_k on antenna 0 (demod_impl.cc:380-392, demod2_impl.cc:356-371); MU: from both LTFs with
         * this user's row of the P matrix, pilots of the second LTF sign-flipped (LTF_NL_28_F_FLOAT_VHT22) */
#pragma unroll
        for (int h = 0; h < 2; h++) {
            int i = lane + 32 * h;
            float l = T->ltf_nl[i];
            float2 v = make_float2(0.f, 0.f);
            if (l != 0.0f) {
                if (!mu) {
                    v = make_float2(__fdiv_rn(S.F[2][i].x, l), __fdiv_rn(S.F[2][i].y, l));
                } else {
                    const cf32 f1 = c_make(S.F[2][i].x, S.F[2][i].y), f2 = c_make(S.F[1][i].x, S.F[1][i].y);
                    cf32 hv;
                    if (mu_pos == 0) {
                        hv = c_divf(c_sub(f1, f2), __fmul_rn(l, 2.0f));
                    } else {
                        const float l22 = (i == 7 || i == 21 || i == 43 || i == 57) ? -l : l;
                        hv = c_divf(c_add(c_divf(f1, l), c_divf(f2, l22)), 2.0f);
                    }
                    v = make_float2(hv.x, hv.y);
                }
            }
            cho[i] = v;
            S.eq[1][i] = v; /* keep H for SIG-B */
        }
        __syncwarp();
        if (fmt == W80211_F_VHT) {
            __syncwarp();
#pragma unroll
            for (int h = 0; h < 2; h++) {
                int i = lane + 32 * h;
                if (T->ltf_nl[i] != 0.0f) {
                    cf32 e = c_div(c_make(S.F[0][i].x, S.F[0][i].y), c_make(S.eq[1][i].x, S.eq[1][i].y));
                    S.eq[0][i] = make_float2(e.x, e.y);
                }
            }
            __syncwarp();
            cf32 psb = c_conj(c_add(c_add(c_sub(c_make(S.eq[0][7].x, S.eq[0][7].y), c_make(S.eq[0][21].x, S.eq[0][21].y)),
                                          c_make(S.eq[0][43].x, S.eq[0][43].y)),
                                    c_make(S.eq[0][57].x, S.eq[0][57].y)));
            float pab = c_abs(psb);
#pragma unroll
            for (int h = 0; h < 2; h++) {
                int i = lane + 32 * h;
                int j = T->bin_j_nl[i];
                if (j >= 0) {
                    cf32 q = c_divf(c_mul(c_make(S.eq[0][i].x, S.eq[0][i].y), psb), pab);
                    S.qam[0][j] = make_float2(q.x, q.y);
                    S.llr[0][T->sigb_deint[j]] = q.x;
                }
            }
            __syncwarp();
        }
    } else {
        /* 2x2: second LTF of both antennas, then H, (H H^H)^-1 and the pilot references
         * (demod2_impl.cc:372-464) */
        float wr[8], wi[8], re[8], im[8];
        fft64_twiddles(g, wr, wi);
        {
            const float2* src = grp == 0 ? x0 + 328 : (grp == 1 ? x1 + 328 : (grp == 2 ? x0 + sigb_at : x1 + sigb_at));
            bool valid = grp < 2 || fmt == W80211_F_VHT;
            fft64_lane8(src, g, S.xch[grp], re, im, wr, wi, valid);
            /* F[2],F[3] hold LTF1 ant0/ant1; put LTF2 into F[0],F[1] and SIG-B into eq[] scratch via registers */
        }
        float2 sb[8];
#pragma unroll
        for (int k2 = 0; k2 < 8; k2++) sb[k2] = make_float2(re[k2], im[k2]);
        if (grp < 2) {
#pragma unroll
            for (int k2 = 0; k2 < 8; k2++) S.F[grp][g + 8 * k2] = sb[k2];
        }
        __syncwarp();
        float2* H = cho;             /* [64][4] */
        float2* HI = cho + 256;      /* [64][4] */
        float2* PR = cho + 512;      /* [2][4] pilot references d_pilotNlLtf / d_pilotNlLtf2 */
#pragma unroll
        for (int h = 0; h < 2; h++) {
            int i = lane + 32 * h;
            float l2 = __fmul_rn(T->ltf_nl[i], 0.5f); /* reference LTF_NL_28_F_FLOAT2 */
            if (l2 != 0.0f) {
                cf32 a1 = c_make(S.F[2][i].x, S.F[2][i].y), a2 = c_make(S.F[3][i].x, S.F[3][i].y);
                cf32 b1 = c_make(S.F[0][i].x, S.F[0][i].y), b2 = c_make(S.F[1][i].x, S.F[1][i].y);
                cf32 h0 = c_scale(c_sub(a1, b1), l2), h1 = c_scale(c_sub(a2, b2), l2);
                cf32 h2 = c_scale(c_add(a1, b1), l2), h3 = c_scale(c_add(a2, b2), l2);
                H[i * 4 + 0] = make_float2(h0.x, h0.y);
                H[i * 4 + 1] = make_float2(h1.x, h1.y);
                H[i * 4 + 2] = make_float2(h2.x, h2.y);
                H[i * 4 + 3] = make_float2(h3.x, h3.y);
            } else {
                for (int q = 0; q < 4; q++) H[i * 4 + q] = make_float2(0.f, 0.f);
            }
        }
        __syncwarp();
        if (fmt == W80211_F_VHT && lane < 16) {
            /* VHT-LTF pilots are not P-mapped: replace the four pilot bins by the mean of their neighbours */
            const int pb[4] = {7, 21, 43, 57};
            int i = pb[lane >> 2], q = lane & 3;
            cf32 lo = c_make(H[(i - 1) * 4 + q].x, H[(i - 1) * 4 + q].y), hi = c_make(H[(i + 1) * 4 + q].x, H[(i + 1) * 4 + q].y);
            cf32 v = c_divf(c_add(lo, hi), 2.0f);
            H[i * 4 + q] = make_float2(v.x, v.y);
        }
        __syncwarp();
#pragma unroll
        for (int h = 0; h < 2; h++) {
            int i = lane + 32 * h;
            if (T->ltf_nl[i] != 0.0f) {
                cf32 h0 = c_make(H[i * 4].x, H[i * 4].y), h1 = c_make(H[i * 4 + 1].x, H[i * 4 + 1].y);
                cf32 h2 = c_make(H[i * 4 + 2].x, H[i * 4 + 2].y), h3 = c_make(H[i * 4 + 3].x, H[i * 4 + 3].y);
                cf32 a = c_add(c_mul(h0, c_conj(h0)), c_mul(h1, c_conj(h1)));
                cf32 bb = c_add(c_mul(h0, c_conj(h2)), c_mul(h1, c_conj(h3)));
                cf32 c = c_add(c_mul(h2, c_conj(h0)), c_mul(h3, c_conj(h1)));
                cf32 dd = c_add(c_mul(h2, c_conj(h2)), c_mul(h3, c_conj(h3)));
                cf32 inv = c_div(c_make(1.0f, 0.0f), c_sub(c_mul(a, dd), c_mul(bb, c)));
                cf32 i0 = c_mul(inv, dd), i1 = c_mul(c_neg(inv), bb), i2 = c_mul(c_neg(inv), c), i3 = c_mul(inv, a);
                HI[i * 4 + 0] = make_float2(i0.x, i0.y);
                HI[i * 4 + 1] = make_float2(i1.x, i1.y);
                HI[i * 4 + 2] = make_float2(i2.x, i2.y);
                HI[i * 4 + 3] = make_float2(i3.x, i3.y);
            } else {
                for (int q = 0; q < 4; q++) HI[i * 4 + q] = make_float2(0.f, 0.f);
            }
        }
        __syncwarp();
        if (lane < 4) {
            /* pilot references from LTF-1 through the equaliser; order -21,-7,+7,+21 -> bins 43,57,7,21;
             * the k=57 entry carries a minus sign (demod2_impl.cc:430-463) */
            const int pb[4] = {43, 57, 7, 21};
            int i = pb[lane];
            cf32 y1 = c_make(S.F[2][i].x, S.F[2][i].y), y2 = c_make(S.F[3][i].x, S.F[3][i].y);
            cf32 h0 = c_make(H[i * 4].x, H[i * 4].y), h1 = c_make(H[i * 4 + 1].x, H[i * 4 + 1].y);
            cf32 h2 = c_make(H[i * 4 + 2].x, H[i * 4 + 2].y), h3 = c_make(H[i * 4 + 3].x, H[i * 4 + 3].y);
            cf32 t1 = c_add(c_mul(y1, c_conj(h0)), c_mul(y2, c_conj(h1)));
            cf32 t2 = c_add(c_mul(y1, c_conj(h2)), c_mul(y2, c_conj(h3)));
            cf32 s1 = c_add(c_mul(t1, c_make(HI[i * 4].x, HI[i * 4].y)), c_mul(t2, c_make(HI[i * 4 + 2].x, HI[i * 4 + 2].y)));
            cf32 s2 = c_add(c_mul(t1, c_make(HI[i * 4 + 1].x, HI[i * 4 + 1].y)), c_mul(t2, c_make(HI[i * 4 + 3].x, HI[i * 4 + 3].y)));
            if (lane == 1) {
                s1 = c_neg(s1);
                s2 = c_neg(s2);
            }
            s1 = c_conj(s1);
            s2 = c_conj(s2);
            PR[lane] = make_float2(s1.x, s1.y);
            PR[4 + lane] = make_float2(s2.x, s2.y);
        }
        __syncwarp();
        if (fmt == W80211_F_VHT) {
            /* VHT-SIG-B, two streams (demod2_impl.cc:663-707) */
            if (grp >= 2) {
#pragma unroll
                for (int k2 = 0; k2 < 8; k2++) S.F[grp - 2][g + 8 * k2] = sb[k2];
            }
            __syncwarp();
#pragma unroll
            for (int h = 0; h < 2; h++) {
                int i = lane + 32 * h;
                if (T->ltf_nl[i] != 0.0f) {
                    cf32 y1 = c_make(S.F[0][i].x, S.F[0][i].y), y2 = c_make(S.F[1][i].x, S.F[1][i].y);
                    cf32 h0 = c_make(H[i * 4].x, H[i * 4].y), h1 = c_make(H[i * 4 + 1].x, H[i * 4 + 1].y);
                    cf32 h2 = c_make(H[i * 4 + 2].x, H[i * 4 + 2].y), h3 = c_make(H[i * 4 + 3].x, H[i * 4 + 3].y);
                    cf32 t1 = c_add(c_mul(y1, c_conj(h0)), c_mul(y2, c_conj(h1)));
                    cf32 t2 = c_add(c_mul(y1, c_conj(h2)), c_mul(y2, c_conj(h3)));
                    cf32 s1 = c_add(c_mul(t1, c_make(HI[i * 4].x, HI[i * 4].y)), c_mul(t2, c_make(HI[i * 4 + 2].x, HI[i * 4 + 2].y)));
                    cf32 s2 = c_add(c_mul(t1, c_make(HI[i * 4 + 1].x, HI[i * 4 + 1].y)), c_mul(t2, c_make(HI[i * 4 + 3].x, HI[i * 4 + 3].y)));
                    S.eq[0][i] = make_float2(s1.x, s1.y);
                    S.eq[1][i] = make_float2(s2.x, s2.y);
                }
            }
            __syncwarp();
            auto E = [&](int s, int i) { return c_make(S.eq[s][i].x, S.eq[s][i].y); };
            auto P = [&](int s, int k) { return c_make(PR[4 * s + k].x, PR[4 * s + k].y); };
            cf32 sum = c_mul(E(0, 7), P(0, 2));
            sum = c_sub(sum, c_mul(E(0, 21), P(0, 3)));
            sum = c_add(sum, c_mul(E(0, 43), P(0, 0)));
            sum = c_add(sum, c_mul(E(0, 57), P(0, 1)));
            sum = c_add(sum, c_mul(E(1, 7), P(1, 2)));
            sum = c_sub(sum, c_mul(E(1, 21), P(1, 3)));
            sum = c_add(sum, c_mul(E(1, 43), P(1, 0)));
            sum = c_add(sum, c_mul(E(1, 57), P(1, 1)));
            cf32 psb = c_conj(sum);
            float pab = c_abs(psb);
#pragma unroll
            for (int h = 0; h < 2; h++) {
                int i = lane + 32 * h;
                int j = T->bin_j_nl[i];
                if (j >= 0) {
                    cf32 q0 = c_divf(c_mul(E(0, i), psb), pab);
                    cf32 q1 = c_divf(c_mul(E(1, i), psb), pab);
                    S.qam[0][j] = make_float2(q0.x, q0.y);
                    S.qam[1][j] = make_float2(q1.x, q1.y);
                    S.llr[0][T->sigb_deint[j]] = __fdiv_rn(__fadd_rn(q0.x, q1.x), 2.0f);
                }
            }
            __syncwarp();
        }
    }

    if (fmt == W80211_F_VHT) {
        /* SIG-B: 26-step Viterbi, parse, per-stream SNR from the re-encoded symbol */
        sig_viterbi(S, S.llr[0], 26, lane);
        /* bccEncoder (cloud80211phy.cc:2622-2644) + procIntelVhtB20 */
        if (lane < 26) {
            int st = 0;
            for (int k = 0; k < 6; k++) {
                int idx = lane - 1 - k;
                st |= (idx >= 0 ? b[idx] : 0) << (5 - k);
            }
            int lab = w_branch_label(st, b[lane]);
            S.coded[T->sigb_int[2 * lane]] = (lab >> 1) & 1;
            S.coded[T->sigb_int[2 * lane + 1]] = lab & 1;
        }
        __syncwarp();
        if (lane < (nant == 1 ? 1 : d.nss)) {
            double np = 0.0;
            for (int i = 0; i < 52; i++) {
                float qr = __fsub_rn(S.qam[lane][i].x, S.coded[i] ? 1.0f : -1.0f);
                float qi = __fsub_rn(S.qam[lane][i].y, 0.0f);
                np += (double)__fadd_rn(__fmul_rn(qr, qr), __fmul_rn(qi, qi));
            }
            float ss = (float)(log10(52.0 / np) * 10.0);
            if (lane == 0) S.llr[1][0] = ss; else S.llr[1][1] = ss;
        }
        __syncwarp();
        d.sssnr0 = S.llr[1][0];
        d.sssnr1 = (nant == 2 && d.nss == 2) ? S.llr[1][1] : 0.0f;
        /* signalParserVhtB (cloud80211phy.cc:1180-1223): MU carries a 16-bit length and the user's MCS */
        if (mu) {
            d.len = bits_le(b, 0, 16) * 4;
            d.mcs = bits_le(b, 16, 4);
            if (d.mcs > 9) {
                d.status = W80211_ST_UNSUPPORTED;
                if (lane == 0) desc_out[p] = d;
                return;
            }
            set_mod_nl(d, d.mcs);
            d.nltf = 2;
            d.nsym = (d.len * 8 + 22 + d.ndbps - 1) / d.ndbps;
        } else if (b[17] + b[18] + b[19] == 3) {
            d.len = bits_le(b, 0, 17) * 4;
            d.nsym = (d.len * 8 + 22 + d.ndbps - 1) / d.ndbps;
        } else {
            d.len = bits_le(b, 0, 20) == 0x422E0 ? 0 : -1; /* 0b01000010001011100000: the NDP SIG-B pattern */
            d.nsym = d.len;
        }
        d.trellis = d.nsym * d.ndbps;
    }

    /* gates (demod_impl.cc:159,186; demod2_impl.cc:159,186) */
    const int nlsym = (pk.l_len * 8 + 22 + 23) / 24;
    bool len_ok = (fmt == W80211_F_VHT && nant == 1) ? (d.len >= 0) : (d.len > 0);
    len_ok = len_ok && d.len <= 4095;
    long long need = (long long)d.nsym * d.nsymsamp + hdr_len;
    if (!(len_ok && (long long)nlsym * 80 >= need)) {
        d.status = W80211_ST_DROP_GATE;
    } else if (d.nsym == 0) {
        d.status = W80211_ST_NDP;
    } else if (sounding) {
        d.status = W80211_ST_UNSUPPORTED; /* two streams of data on one antenna: the reference's output is undefined here */
    } else if (need > (long long)pk.nsamp) {
        d.status = W80211_ST_BAD_INPUT; /* nsamp tag shorter than the L-SIG duration it came from */
    }
    d.data_off = hdr_len;
    d.total = d.nsym > 0 ? d.nsym * d.ncbps : 0;
    if (lane == 0) desc_out[p] = d;
}

void w_launch_header(cudaStream_t st, const W80211Tables* tab, int npkt, int nant, const float2* iq0, const float2* iq1,
                     unsigned long long nsamples, const w80211_pkt_in* pk, const float2* chan_in, w80211_desc* desc,
                     float2* chan_out, int chan_stride, int mu_pos)
{
    if (npkt <= 0) return;
    hdr_kernel<<<npkt, 32, 0, st>>>(tab, npkt, nant, iq0, iq1, nsamples, pk, chan_in, desc, chan_out, chan_stride, mu_pos);
}

/* ------------------------------------------------------------------------------------------------
 * L-SIG of one sync flag, one warp per flag: the S_DEMOD state of the reference's signal / signal2
 * blocks (lib/signal_impl.cc:108-160, lib/signal2_impl.cc:109-163).
 *   rotate the two L-LTF symbols and the L-SIG symbol (offsets +8, +72, +152 from the flag) by
 *   exp(j*(i+offset)*rad), FFT each, H = (LTF1 + LTF2) / (2*L) on the 52 legacy bins,
 *   procLHSigDemodDeint (cloud80211phy.cc:609-627), 24-step Viterbi (:2001-2088),
 *   signalCheckLegacy (:650-728).
 * res[p] = (1, rate index, LENGTH, nsym*80) when the check passes, (0,..) when it fails, (-1,..) when
 * the flag has no tags or fewer than 224 samples follow it.  chan_all[p] = the "chan" tag: H on the
 * legacy bins, zero elsewhere (d_h starts as zeros and only those bins are ever written). */
__global__ void __launch_bounds__(32) lsig_kernel(const W80211Tables* __restrict__ T, int nsync,
                                                  const w80211_sync* __restrict__ syncs, const float2* __restrict__ iq0,
                                                  unsigned long long nsamples, int4* __restrict__ res,
                                                  float2* __restrict__ chan_all)
{
    __shared__ HdrSmem S;
    const int p = blockIdx.x;
    if (p >= nsync) return;
    const int lane = threadIdx.x;
    const int grp = lane >> 3, g = lane & 7;
    const w80211_sync sy = syncs[p];
    float2* cho = chan_all + (size_t)p * 64;
    if (!(sy.flags & W80211_SYNC_TAGGED) || sy.pos + 224ull > nsamples) {
        if (lane == 0) res[p] = make_int4(-1, 0, 0, 0);
        cho[lane] = make_float2(0.f, 0.f);
        cho[lane + 32] = make_float2(0.f, 0.f);
        return;
    }
    const float2* x = iq0 + sy.pos;
    {
        float wr[8], wi[8], re[8], im[8];
        fft64_twiddles(g, wr, wi);
        const int base = grp == 0 ? 8 : (grp == 1 ? 72 : 152);
#pragma unroll
        for (int n1 = 0; n1 < 8; n1++) {
            re[n1] = 0.f;
            im[n1] = 0.f;
            if (grp < 3) {
                const int k = base + 8 * n1 + g;
                const float2 v = __ldg(x + k);
                float sn, cs;
                sincosf(__fmul_rn((float)k, sy.rad), &sn, &cs);
                const cf32 r = c_mul(c_make(v.x, v.y), c_make(cs, sn));
                re[n1] = r.x;
                im[n1] = r.y;
            }
        }
        fft64_lane8_regs(g, S.xch[grp], re, im, wr, wi);
#pragma unroll
        for (int k2 = 0; k2 < 8; k2++) S.F[grp][g + 8 * k2] = make_float2(re[k2], im[k2]);
    }
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int i = lane + 32 * h;
        const bool used = T->sig_demap[i] >= 0 || i == 7 || i == 21 || i == 43 || i == 57;
        cf32 hh = c_make(0.f, 0.f);
        if (used) {
            /* LTF_L_26_F_FLOAT2[i] = 2 * L-LTF sign, the HT-LTF sign on the legacy bins */
            const float l2 = __fmul_rn(2.0f, T->ltf_nl[i]);
            hh = c_divf(c_add(c_make(S.F[0][i].x, S.F[0][i].y), c_make(S.F[1][i].x, S.F[1][i].y)), l2);
        }
        S.eq[0][i] = make_float2(hh.x, hh.y);
        cho[i] = make_float2(hh.x, hh.y);
    }
    __syncwarp();
    const float2* hl = S.eq[0];
    const cf32 e7 = c_div(c_make(S.F[2][7].x, S.F[2][7].y), c_make(hl[7].x, hl[7].y));
    const cf32 e21 = c_div(c_make(S.F[2][21].x, S.F[2][21].y), c_make(hl[21].x, hl[21].y));
    const cf32 e43 = c_div(c_make(S.F[2][43].x, S.F[2][43].y), c_make(hl[43].x, hl[43].y));
    const cf32 e57 = c_div(c_make(S.F[2][57].x, S.F[2][57].y), c_make(hl[57].x, hl[57].y));
    const cf32 ps = c_conj(c_add(c_add(c_sub(e7, e21), e43), e57));
    const float pa = c_abs(ps);
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int i = lane + 32 * h;
        const int dm = T->sig_demap[i];
        if (dm >= 0) {
            const cf32 m = c_divf(c_mul(c_div(c_make(S.F[2][i].x, S.F[2][i].y), c_make(hl[i].x, hl[i].y)), ps), pa);
            S.llr[0][dm] = m.x;
        }
    }
    __syncwarp();
    sig_viterbi(S, S.llr[0], 24, lane);
    if (lane == 0) {
        const uint8_t* b = S.bits;
        int par = 0;
        for (int i = 0; i < 17; i++) par += b[i];
        const int rate = bits_le(b, 0, 4);
        const int len = bits_le(b, 5, 12);
        bool ok = b[3] == 1 && b[4] == 0 && ((par & 1) ^ b[17]) == 0 && len >= 14 && len <= 4095;
        /* RATE field (bit 0 first) -> rate index; unknown patterns fall back to 0 like the reference's default case */
        int mcs = 0;
        switch (rate) {
        case 11: mcs = 0; break;
        case 15: mcs = 1; break;
        case 10: mcs = 2; break;
        case 14: mcs = 3; break;
        case 9: mcs = 4; break;
        case 13: mcs = 5; break;
        case 8: mcs = 6; break;
        case 12: mcs = 7; break;
        default: mcs = 0; break;
        }
        const int nd[8] = {24, 36, 48, 72, 96, 144, 192, 216};
        const int nsym = (len * 8 + 22 + nd[mcs] - 1) / nd[mcs];
        res[p] = ok ? make_int4(1, mcs, len, nsym * 80) : make_int4(0, 0, 0, 0);
    }
}

void w_launch_lsig(cudaStream_t st, const W80211Tables* tab, int nsync, const w80211_sync* syncs, const float2* iq0,
                   unsigned long long nsamples, int4* res, float2* chan_all)
{
    if (nsync <= 0) return;
    lsig_kernel<<<nsync, 32, 0, st>>>(tab, nsync, syncs, iq0, nsamples, res, chan_all);
}
