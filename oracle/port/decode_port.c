/* decode_port.c -- plain-C restatement of the reference's decode block, for checking only.
 *
 * TEST INFRASTRUCTURE.  Not part of the product: nothing under fork-gr-ieee80211_b200/, gr/ or include/
 * links, loads or calls it.  Used by tests/ (second opinion beside oracle/_ref, and the only CPU checker of the
 * decode half on a machine where the reference tree is absent) and as bench.py's `cpu_baseline` of kind "port"
 * when oracle/_ref is missing.
 *
 * What it restates (reference = /root/reference/lib/decode_impl.cc, cloud80211phy.cc):
 *   port_viterbi     vstb_init + vstb_update + vstb_end   decode_impl.cc:164-302
 *                    (depuncture patterns SV_PUNC_* cloud80211phy.cc:1857-1860, trellis tables :1864-1887)
 *   port_descramble  descramble                            decode_impl.cc:304-323
 *   port_assemble    packetAssemble + boost::crc_32_type   decode_impl.cc:325-520, decode_impl.h:84
 * Pinned by tests/test_port.py against the reference's golden fixture (tests/golden/fixture_*.npz: LLR stream ->
 * the 94-byte frame of tools/debug.txt:5) and against oracle/_ref on random packets of every code rate.
 * The per-symbol demodulation is restated next door, in demod_port.c; the header stage (SIG parsing) is checked by
 * oracle/_ref only (the reference sources themselves).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define T_MAX 32782 /* DECODE_T_MAX, decode_impl.h:36 */
#define B_MAX 4095  /* DECODE_B_MAX, decode_impl.h:35 */

/* K = 7, generators 133 / 171 (octal); state = the six previous input bits, newest in bit 5.
 * next state and 2-bit output label (A << 1 | B) of input bit b: what SV_STATE_NEXT / SV_STATE_OUTPUT tabulate */
static int trellis_next(int s, int b) { return (s >> 1) | (b << 5); }
static int trellis_label(int s, int b)
{
    int x1 = (s >> 5) & 1, x2 = (s >> 4) & 1, x3 = (s >> 3) & 1, x5 = (s >> 1) & 1, x6 = s & 1;
    return ((b ^ x2 ^ x3 ^ x5 ^ x6) << 1) | (b ^ x1 ^ x2 ^ x3 ^ x6);
}
void port_trellis_tables(int* next128, int* label128)
{
    for (int s = 0; s < 64; s++)
        for (int b = 0; b < 2; b++) {
            next128[2 * s + b] = trellis_next(s, b);
            label128[2 * s + b] = trellis_label(s, b);
        }
}

/* Soft Viterbi with the reference's exact arithmetic: correlation metric in binary32, sentinel -1e15, candidates
 * visited in ascending source order with a strict '>' (so the lower source wins ties), survivor of a never-updated
 * state = 0, traceback over the whole trellis from state 0.  Returns LLRs consumed, or -1. */
int port_viterbi(const float* llr, int total, int trellis, int cr, uint8_t* bits /* [trellis] */)
{
    static const int punc[4][10] = { { 1, 1 }, { 1, 1, 1, 0 }, { 1, 1, 1, 0, 0, 1 }, { 1, 1, 1, 0, 0, 1, 1, 0, 0, 1 } };
    static const int plen[4] = { 2, 4, 6, 10 };
    if (cr < 0 || cr > 3 || trellis <= 0 || trellis > T_MAX) return -1;
    uint8_t* his = (uint8_t*)calloc((size_t)(trellis + 1) * 64, 1); /* his[t][state] = winning source */
    if (!his) return -1;
    static int NEXT[64][2], LABEL[64][2], ready = 0;
    if (!ready) {
        for (int s = 0; s < 64; s++)
            for (int b = 0; b < 2; b++) {
                NEXT[s][b] = trellis_next(s, b);
                LABEL[s][b] = trellis_label(s, b);
            }
        ready = 1;
    }
    float a0[64], a1[64], *pre = a0, *cur = a1;
    for (int i = 0; i < 64; i++) a0[i] = a1[i] = -1000000000000000.0f;
    a0[0] = 0.0f;
    int used = 0, pp = 0;
    for (int t = 0; t < trellis; t++) {
        if (used + punc[cr][pp] + punc[cr][pp + 1] > total) { free(his); return -1; }
        const float t0 = punc[cr][pp] ? llr[used++] : 0.0f;
        const float t1 = punc[cr][pp + 1] ? llr[used++] : 0.0f;
        float tab[4];
        tab[0] = 0.0f;
        tab[1] = t1;
        tab[2] = t0;
        tab[3] = t1 + t0;
        uint8_t* h = his + (size_t)(t + 1) * 64;
        for (int s = 0; s < 64; s++)
            for (int b = 0; b < 2; b++) {
                float cand = pre[s] + tab[LABEL[s][b]];
                int n = NEXT[s][b];
                if (cand > cur[n]) {
                    cur[n] = cand;
                    h[n] = (uint8_t)s;
                }
            }
        float* sw = pre;
        pre = cur;
        cur = sw;
        for (int i = 0; i < 64; i++) cur[i] = -1000000000000000.0f;
        pp += 2;
        if (pp >= plen[cr]) pp = 0;
    }
    int st = 0;
    for (int t = trellis; t > 0; t--) {
        int prev = his[(size_t)t * 64 + st];
        bits[t - 1] = (uint8_t)(st == trellis_next(prev, 1));
        st = prev;
    }
    free(his);
    return used;
}

/* x^7 + x^4 + 1 scrambler: state from the first seven bits, which are then forced to 0 */
void port_descramble(const uint8_t* in, uint8_t* out, int trellis)
{
    int state = 0;
    for (int i = 0; i < 7 && i < trellis; i++)
        if (in[i]) state |= 1 << (6 - i);
    for (int i = 0; i < 7 && i < trellis; i++) out[i] = 0;
    for (int i = 7; i < trellis; i++) {
        int fb = ((state >> 6) & 1) ^ ((state >> 3) & 1);
        out[i] = (uint8_t)(fb ^ (in[i] & 1));
        state = ((state << 1) & 0x7e) | fb;
    }
}

static uint32_t crc32_bytes(const uint8_t* p, int n)
{
    static uint32_t tab[256];
    static int ready = 0;
    if (!ready) {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            tab[i] = c;
        }
        ready = 1;
    }
    uint32_t c = 0xFFFFFFFFu;
    for (int i = 0; i < n; i++) c = tab[(c ^ p[i]) & 0xff] ^ (c >> 8);
    return c ^ 0xFFFFFFFFu;
}

static void pack_bytes(const uint8_t* bit, int nbytes, uint8_t* dst)
{
    for (int i = 0; i < nbytes; i++) {
        int v = 0;
        for (int j = 0; j < 8; j++) v |= (bit[i * 8 + j] & 1) << j;
        dst[i] = (uint8_t)v;
    }
}

/* Messages the reference publishes for one decoded packet: [format, len lo, len hi, MPDU..., mcs] for every MPDU
 * whose CRC-32 over MPDU + FCS leaves the residue 558161692.  msgs: concatenation of (uint16 length, bytes).
 * *nfail counts MPDUs walked whose FCS is wrong.  Returns the number of messages. */
int port_assemble(const uint8_t* ubits, int format, int len, int mcs, int ampdu, int trellis, uint8_t* msgs, int cap, int* nbytes,
                  int* nfail)
{
    int nmsg = 0, used = 0, fail = 0;
    static uint8_t pkt[B_MAX + 8];
    if (format == 2) { /* VHT: walk the A-MPDU delimiters, quirks of the reference kept (no delimiter CRC / signature check,
                          the length accumulates by OR from one subframe to the next) */
        int procd = 16;
        if (procd < trellis) {
            const uint8_t* p = ubits + 16;
            int mlen = 0;
            for (;;) {
                procd += 32;
                if (procd > trellis) break;
                int eof = p[0];
                mlen |= p[2] << 12;
                mlen |= p[3] << 13;
                for (int i = 0; i < 12; i++) mlen |= p[4 + i] << i;
                int padded = (mlen / 4 + ((mlen % 4) != 0)) * 4;
                procd += padded * 8;
                if (procd > trellis) break;
                p += 32;
                pkt[0] = (uint8_t)format;
                pkt[1] = (uint8_t)(mlen % 256);
                pkt[2] = (uint8_t)(mlen / 256);
                pack_bytes(p, mlen, pkt + 3);
                p += padded * 8;
                if (crc32_bytes(pkt + 3, mlen) == 558161692u) {
                    pkt[mlen + 3] = (uint8_t)mcs;
                    mlen += 4;
                    if (used + 2 + mlen <= cap) {
                        msgs[used] = (uint8_t)(mlen & 0xff);
                        msgs[used + 1] = (uint8_t)(mlen >> 8);
                        memcpy(msgs + used + 2, pkt, (size_t)mlen);
                        used += 2 + mlen;
                        nmsg++;
                    }
                } else {
                    fail++;
                }
                if (eof) break;
            }
        }
    } else if (!ampdu) { /* legacy and HT: one MPDU of `len` bytes after the 16 SERVICE bits; HT A-MPDU is a stub in the reference */
        if (len >= 0 && len <= B_MAX && 16 + 8 * len <= trellis) {
            pkt[0] = (uint8_t)format;
            pkt[1] = (uint8_t)(len % 256);
            pkt[2] = (uint8_t)(len / 256);
            pack_bytes(ubits + 16, len, pkt + 3);
            if (crc32_bytes(pkt + 3, len) == 558161692u) {
                pkt[3 + len] = (uint8_t)mcs;
                int n = len + 4;
                if (2 + n <= cap) {
                    msgs[0] = (uint8_t)(n & 0xff);
                    msgs[1] = (uint8_t)(n >> 8);
                    memcpy(msgs + 2, pkt, (size_t)n);
                    used = 2 + n;
                    nmsg = 1;
                }
            } else {
                fail++;
            }
        }
    }
    *nbytes = used;
    if (nfail) *nfail = fail;
    return nmsg;
}

/* One packet of the decode block's input (tags + `total` LLRs) -> messages.  Gates as at decode_impl.cc:93. */
int port_decode_packet(const float* llr, int total, int format, int mcs, int len, int cr, int ampdu, int trellis, uint8_t* msgs,
                       int cap, int* nbytes, int* nfail)
{
    *nbytes = 0;
    if (nfail) *nfail = 0;
    if (len > B_MAX || trellis > T_MAX || trellis <= 0) return 0;
    uint8_t* sb = (uint8_t*)malloc((size_t)trellis * 2 + 64);
    if (!sb) return -1;
    uint8_t* ub = sb + trellis + 32;
    int r = port_viterbi(llr, total, trellis, cr, sb);
    if (r < 0) {
        free(sb);
        return -1;
    }
    port_descramble(sb, ub, trellis);
    int n = port_assemble(ub, format, len, mcs, ampdu, trellis, msgs, cap, nbytes, nfail);
    free(sb);
    return n;
}
