/* post.cu -- descramble, byte packing, A-MPDU walk and FCS, one warp per decoded packet.
 *
 * Replaces decode_impl::descramble (lib/decode_impl.cc:304-323) and decode_impl::packetAssemble
 * (:325-520) including the boost::crc_32_type check (decode_impl.h:84, :372-374, :451-453).
 *
 *  - descramble: the 7-bit scrambler state is read from the first 7 decoded bits; the 160 bits of the x^7+x^4+1
 *    sequence that follow it come from a table made on the host and are applied 32 bits at a time;
 *  - bytes: LSB first from bit 16, i.e. the descrambled bit words viewed as bytes, offset 2;
 *  - CRC-32 (reflected 0xEDB88320, init/xorout 0xFFFFFFFF): every lane runs a four-bytes-per-step table CRC
 *    (tables made on the host, tables.cc) over its own slice of the MPDU, the 32 partial CRCs are merged with
 *    x^(8n) multiplications in GF(2)[x]/P, and the result is compared with the residue 558161692 the reference
 *    tests for;
 *  - VHT: the delimiter walk of the reference, with its quirks (no delimiter CRC/signature check,
 *    length bits OR-ed into a value that survives from the previous subframe, +4 after a pass).
 */
#include "common.cuh"

#ifndef POST_WARPS
#define POST_WARPS 8 /* one CRC table copy per 8 packets; 37 KB of static shared memory, 6 blocks per SM */
#endif
#define POST_MAXW 1028 /* ceil(32782/32) = 1025 words, padded */
#define CRC_POLY 0xEDB88320u
#define CRC_RESIDUE 558161692u

struct PostSmem {
    uint32_t crc_tab[4][256];
    uint32_t x2n[32]; /* x^(2^k) mod P, reflected */
    uint32_t dw[POST_WARPS][POST_MAXW];
    uint32_t seq[POST_WARPS][8];
    const uint32_t* src[POST_WARPS]; /* word 0 of each packet of the block in the traceback's output */
    int T[POST_WARPS];
};
static_assert(POST_MAXW % 32 == 4 && (POST_WARPS & (POST_WARPS - 1)) == 0 && POST_WARPS <= 32, "transposing store: row stride 4 mod 32 banks");

/* a(x) * b(x) mod P, reflected bit order (bit 31 = x^0) */
__device__ __forceinline__ uint32_t gf_mul(uint32_t a, uint32_t b)
{
    uint32_t p = 0;
#pragma unroll 4
    for (int i = 31; i >= 0; i--) {
        if ((a >> i) & 1u) p ^= b;
        b = (b & 1u) ? (b >> 1) ^ CRC_POLY : (b >> 1);
    }
    return p;
}

/* x^(8*nbytes) mod P */
__device__ __forceinline__ uint32_t x_pow_bytes(const uint32_t* x2n, uint32_t nbytes)
{
    uint32_t p = 0x80000000u; /* x^0 */
    int k = 3;
    while (nbytes) {
        if (nbytes & 1u) p = gf_mul(x2n[k & 31], p);
        nbytes >>= 1;
        k++;
    }
    return p;
}

/* standard CRC-32 of the n bytes that start at byte 2 of dw[woff] (PSDU bytes sit at bit offset 16 of the descrambled
 * words; A-MPDU subframes start on 4-byte boundaries), computed by the whole warp; every lane gets the result */
__device__ uint32_t warp_crc32(const PostSmem& S, const uint32_t* __restrict__ xpow8, const uint32_t* dw, int woff, int n, int lane)
{
    if (n <= 0) return 0u;
    int c = (n + 31) >> 5;
    c = (c + 3) & ~3;
    if (!((c >> 2) & 1)) c += 4; /* odd number of words per lane: conflict-free shared reads */
    int beg = lane * c, end = min(n, beg + c);
    uint32_t crc = 0u;
    if (beg < end) {
        crc = 0xFFFFFFFFu;
        const uint32_t* w = dw + woff + (beg >> 2);
        uint32_t lo = w[0];
        int k = beg;
        for (; k + 4 <= end; k += 4) {
            const uint32_t hi = *++w;
            const uint32_t x = crc ^ __funnelshift_r(lo, hi, 16);
            lo = hi;
            crc = S.crc_tab[3][x & 0xffu] ^ S.crc_tab[2][(x >> 8) & 0xffu] ^ S.crc_tab[1][(x >> 16) & 0xffu] ^ S.crc_tab[0][x >> 24];
        }
        const uint8_t* by = reinterpret_cast<const uint8_t*>(dw + woff) + 2;
        for (; k < end; k++) crc = S.crc_tab[0][(crc ^ by[k]) & 0xffu] ^ (crc >> 8);
        crc ^= 0xFFFFFFFFu;
        int after = n - end;
        if (after > 0) crc = gf_mul(after < W_XPOW8 ? __ldg(&xpow8[after]) : x_pow_bytes(S.x2n, (uint32_t)after), crc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) crc ^= __shfl_xor_sync(0xffffffffu, crc, o);
    return crc;
}

__device__ __forceinline__ void emit(int lane, uint32_t pkt, uint32_t off, uint32_t len, bool ok, w80211_mpdu* mpdus,
                                     unsigned int cap, unsigned int* counters)
{
    if (lane == 0) {
        unsigned int idx = atomicAdd(&counters[0], 1u);
        if (idx < cap) {
            w80211_mpdu r;
            r.pkt = pkt;
            r.off = off;
            r.len = len;
            r.crc_ok = ok ? 1u : 0u;
            mpdus[idx] = r;
        }
        if (ok) {
            atomicAdd(&counters[1], 1u);
            atomicAdd(reinterpret_cast<unsigned long long*>(&counters[2]), (unsigned long long)len * 8ull);
        }
    }
}

/* counters: [0] MPDU records, [1] FCS-valid MPDUs, [2..3] 64-bit sum of 8*len over FCS-valid MPDUs */
__global__ void __launch_bounds__(POST_WARPS * 32) post_kernel(const W80211Tables* __restrict__ tab, int ndec, const int4* __restrict__ grp, const int* __restrict__ pkt_slot,
                                                               const int* __restrict__ dec_list, const w80211_desc* __restrict__ desc,
                                                               const uint32_t* __restrict__ bits, const uint32_t* __restrict__ psdu_off,
                                                               uint8_t* __restrict__ out_bytes, w80211_mpdu* __restrict__ mpdus,
                                                               unsigned int mpdu_cap, unsigned int* __restrict__ counters)
{
    __shared__ PostSmem S;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int i = tid; i < 1024; i += blockDim.x) (&S.crc_tab[0][0])[i] = __ldg(&tab->crc_tab[0][0] + i);
    if (tid < 32) S.x2n[tid] = __ldg(&tab->crc_x2n[tid]);
    const int q = blockIdx.x * POST_WARPS + wid;
    const bool valid = q < ndec;
    const int pkt = valid ? dec_list[q] : 0;
    const w80211_desc d = desc[pkt];
    const int T = d.trellis;
    if (valid) {
        const int slot = pkt_slot[pkt];
        const int4 G = grp[slot >> 5];
        const uint32_t* src = bits + (size_t)G.z * 32 + (slot & 31);
        /* scrambler state from the first 7 bits (decode_impl.cc:308-314); the 160 sequence bits of this state come from the
         * host-made table (tables.cc), six words per packet */
        const uint32_t w0 = __ldg(src);
        int state = 0;
        for (int i = 0; i < 7; i++)
            if ((w0 >> i) & 1u) state |= 1 << (6 - i);
        if (lane < 6) S.seq[wid][lane] = __ldg(&tab->scr_seq[state][lane]);
        if (lane == 0) {
            S.src[wid] = src;
            S.T[wid] = T;
        }
    } else if (lane == 0) {
        S.src[wid] = bits;
        S.T[wid] = 0;
    }
    __syncthreads();
    /* The traceback leaves word j of the 32 packets of a group side by side (bits[j][32]); the block's eight packets are
     * eight neighbours there, so the block reads the words together -- a warp takes four rows of eight packets, four whole
     * sectors -- descrambles them on the way and lays them out per packet in shared memory (row stride = 4 mod 32 banks:
     * the transposing store is conflict-free). */
    {
        const int pp = tid & (POST_WARPS - 1), jr = tid / POST_WARPS;
        const int Tp = S.T[pp];
        const int nwp = Tp > 0 ? (Tp + 31) >> 5 : -1;
        const uint32_t* srcp = S.src[pp];
        const uint32_t* seq = S.seq[pp];
        uint32_t* dwp = S.dw[pp];
        for (int j = jr; j < nwp + 1; j += 32) { /* blockDim.x / POST_WARPS rows per pass */
            uint32_t v = 0;
            if (j < nwp) {
                const uint32_t in = __ldg(srcp + (size_t)j * 32);
                if (j == 0) {
                    v = (in ^ (seq[0] << 7)) & ~0x7fu; /* outputs 0..6 are forced to 0 (decode_impl.cc:315) */
                } else {
                    const int o = (32 * j - 7) % 127;
                    const uint32_t sq = __funnelshift_r(seq[o >> 5], seq[(o >> 5) + 1], o & 31);
                    v = in ^ sq;
                }
                const int rem = Tp - 32 * j; /* bits past the trellis are not data */
                if (rem < 32) v &= (1u << rem) - 1u;
            }
            dwp[j] = v;
        }
    }
    __syncthreads();
    if (!valid) return;
    const uint32_t* dw = S.dw[wid];

    /* PSDU bytes: byte k = bits 16+8k .. 23+8k */
    const uint8_t* by = reinterpret_cast<const uint8_t*>(dw) + 2;
    const int navail = T >= 16 ? (T - 16) >> 3 : 0;
    const int nbytes = d.format == W80211_F_VHT ? navail : min(d.len, navail);
    const uint32_t poff = psdu_off[pkt];
    uint32_t* o32 = reinterpret_cast<uint32_t*>(out_bytes + poff);
    for (int k = lane; k < ((nbytes + 3) >> 2); k += 32) o32[k] = (dw[k] >> 16) | (dw[k + 1] << 16);

    if (d.format != W80211_F_VHT) {
        if (d.ampdu) return; /* HT A-MPDU: the reference publishes nothing (decode_impl.cc:432-435) */
        uint32_t crc = warp_crc32(S, tab->crc_xpow8, dw, 0, nbytes, lane);
        emit(lane, (uint32_t)pkt, poff, (uint32_t)nbytes, nbytes == d.len && crc == CRC_RESIDUE, mpdus, mpdu_cap, counters);
        return;
    }
    /* VHT A-MPDU walk (decode_impl.cc:331-426) */
    int procd = 16;
    if (procd >= T) return;
    int pos = 0, tmplen = 0;
    for (;;) {
        procd += 32;
        if (procd > T) break;
        uint32_t b0 = by[pos], b1 = by[pos + 1];
        int eof = b0 & 1u;
        tmplen |= (int)(((b0 >> 2) & 1u) << 12) | (int)(((b0 >> 3) & 1u) << 13) | (int)(b0 >> 4) | (int)(b1 << 4);
        int padded = ((tmplen + 3) >> 2) << 2;
        procd += padded * 8;
        if (procd > T) break;
        pos += 4;
        uint32_t crc = warp_crc32(S, tab->crc_xpow8, dw, pos >> 2, tmplen, lane);
        bool ok = crc == CRC_RESIDUE;
        emit(lane, (uint32_t)pkt, poff + (uint32_t)pos, (uint32_t)tmplen, ok, mpdus, mpdu_cap, counters);
        pos += padded;
        if (ok) tmplen += 4;
        if (eof) break;
    }
}

void w_launch_post(cudaStream_t st, const W80211Tables* tab, int npkt_dec, const int4* grp, const int* pkt_slot, const int* dec_list,
                   const w80211_desc* desc, const uint32_t* bits, const uint32_t* psdu_off, uint8_t* out_bytes,
                   w80211_mpdu* mpdus, unsigned int mpdu_cap, unsigned int* counters)
{
    if (npkt_dec <= 0) return;
    int blocks = (npkt_dec + POST_WARPS - 1) / POST_WARPS;
    post_kernel<<<blocks, POST_WARPS * 32, 0, st>>>(tab, npkt_dec, grp, pkt_slot, dec_list, desc, bits, psdu_off, out_bytes, mpdus,
                                                    mpdu_cap, counters);
}
