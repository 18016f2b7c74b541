/* tables.cc -- constant tables of the receive path, generated from the formulas of
 * IEEE 802.11-2016 (clauses 17, 19, 21) instead of being listed.  tests/test_tables.py checks
 * every one of them against the reference's arrays (cloud80211phy.cc) through oracle/_ref.
 */
#include <cstring>
#include "common.cuh"

extern "C" void w80211_fill_tables(W80211Tables* t);

/* interleaver permutation: coded bit k -> transmitted position (17.3.5.7, 19.3.11.8.3) */
static int interleave_pos(int k, int ncbpss, int nbpscs, bool legacy, int iss)
{
    int s = nbpscs / 2 > 1 ? nbpscs / 2 : 1;
    int ncol = legacy ? 16 : 13;
    int nrow = legacy ? ncbpss / 16 : 4 * nbpscs;
    int i = nrow * (k % ncol) + k / ncol;
    int j = s * (i / s) + (i + ncbpss - (ncol * i) / ncbpss) % s;
    if (!legacy && iss > 1) {
        int rot = ((2 * (iss - 1)) % 3 + 3 * ((iss - 1) / 3)) * 11 * nbpscs;
        j = ((j - rot) % ncbpss + ncbpss) % ncbpss;
    }
    return j;
}

/* a(x) * b(x) mod P in the reflected bit order of CRC-32 (bit 31 = x^0) */
static uint32_t crc_gf_mul(uint32_t a, uint32_t b)
{
    uint32_t p = 0;
    for (int i = 31; i >= 0; i--) {
        if ((a >> i) & 1u) p ^= b;
        b = (b & 1u) ? (b >> 1) ^ 0xEDB88320u : (b >> 1);
    }
    return p;
}

void w80211_fill_tables(W80211Tables* t)
{
    memset(t, 0, sizeof(*t));
    /* CRC-32: byte table, three more tables for four bytes per step, and x^(2^k) mod P for joining partial CRCs */
    for (int i = 0; i < 256; i++) {
        uint32_t c = (uint32_t)i;
        for (int k = 0; k < 8; k++) c = (c & 1u) ? (0xEDB88320u ^ (c >> 1)) : (c >> 1);
        t->crc_tab[0][i] = c;
    }
    for (int s = 1; s < 4; s++)
        for (int i = 0; i < 256; i++) t->crc_tab[s][i] = (t->crc_tab[s - 1][i] >> 8) ^ t->crc_tab[0][t->crc_tab[s - 1][i] & 0xffu];
    t->crc_x2n[0] = 0x40000000u; /* x^1 */
    for (int k = 1; k < 32; k++) t->crc_x2n[k] = crc_gf_mul(t->crc_x2n[k - 1], t->crc_x2n[k - 1]);
    /* descrambler sequences (decode_impl.cc:304-323): state = first seven received bits, bit i at position 6 - i */
    for (int st0 = 0; st0 < 128; st0++) {
        int state = st0;
        for (int k = 0; k < 160; k++) {
            int fb = ((state >> 6) & 1) ^ ((state >> 3) & 1);
            state = ((state << 1) & 0x7e) | fb;
            t->scr_seq[st0][k >> 5] |= (uint32_t)fb << (k & 31);
        }
    }
    t->crc_xpow8[0] = 0x80000000u; /* x^0 */
    for (int n = 1; n < W_XPOW8; n++) t->crc_xpow8[n] = crc_gf_mul(t->crc_x2n[3], t->crc_xpow8[n - 1]);
    /* pilot polarity: scrambler output with the all-ones seed, 0 -> +1, 1 -> -1 (17.3.5.10) */
    int st = 0x7f;
    for (int i = 0; i < 127; i++) {
        int fb = ((st >> 6) & 1) ^ ((st >> 3) & 1);
        st = ((st << 1) & 0x7e) | fb;
        t->pilot_p[i] = fb ? -1.0f : 1.0f;
    }
    t->pilot_p[127] = t->pilot_p[0];

    /* L-LTF, subcarriers -26..26 (17.3.3, Equation 17-8); HT/VHT-LTF adds {1,1} below and {-1,-1} above */
    static const char* lltf = "++--++-+-++++++--++-+-++++0+--++-+-+-----++--+-+-++++";
    int ltf[57];
    ltf[0] = 1;
    ltf[1] = 1;
    for (int i = 0; i < 53; i++) ltf[2 + i] = lltf[i] == '+' ? 1 : (lltf[i] == '-' ? -1 : 0);
    ltf[55] = -1;
    ltf[56] = -1;
    for (int k = -28; k <= 28; k++) t->ltf_nl[(k + 64) % 64] = (float)ltf[k + 28];

    /* bin -> data subcarrier index, ascending frequency, pilots (+-7, +-21) and DC skipped */
    for (int i = 0; i < 64; i++) t->bin_j_l[i] = t->bin_j_nl[i] = t->sig_demap[i] = -1;
    int jl = 0, jn = 0;
    for (int k = -28; k <= 28; k++) {
        if (k == 0 || k == -21 || k == -7 || k == 7 || k == 21) continue;
        int bin = (k + 64) % 64;
        t->bin_j_nl[bin] = (int8_t)jn++;
        if (k >= -26 && k <= 26) t->bin_j_l[bin] = (int8_t)jl++;
    }
    /* SIG symbols: 48 BPSK bits on the legacy tone layout, deinterleaved */
    int de48[48];
    for (int k = 0; k < 48; k++) de48[interleave_pos(k, 48, 1, true, 1)] = k;
    for (int i = 0; i < 64; i++)
        if (t->bin_j_l[i] >= 0) t->sig_demap[i] = (int8_t)de48[t->bin_j_l[i]];
    /* VHT-SIG-B: 52 BPSK bits, HT-style interleaver */
    for (int k = 0; k < 52; k++) {
        int j = interleave_pos(k, 52, 1, false, 1);
        t->sigb_int[k] = (uint8_t)j;
        t->sigb_deint[j] = (uint8_t)k;
    }
    /* data symbols */
    const int nbl[4] = {1, 2, 4, 6}, nbn[5] = {1, 2, 4, 6, 8};
    for (int a = 0; a < 4; a++) {
        int n = 48 * nbl[a];
        for (int k = 0; k < n; k++) t->map[a][interleave_pos(k, n, nbl[a], true, 1)] = (uint16_t)k;
    }
    for (int a = 0; a < 5; a++) {
        int nb = nbn[a], n = 52 * nb, s = nb / 2 > 1 ? nb / 2 : 1;
        for (int k = 0; k < n; k++) {
            t->map[4 + a][interleave_pos(k, n, nb, false, 1)] = (uint16_t)k;
            /* two streams: deinterleave, then the stream de-parser takes s bits from each stream in turn */
            for (int iss = 0; iss < 2; iss++) {
                int pos = (k / s) * 2 * s + iss * s + (k % s);
                t->map[9 + 5 * iss + a][interleave_pos(k, n, nb, false, iss + 1)] = (uint16_t)pos;
            }
        }
    }
    /* the same maps indexed by FFT bin, eight entries per bin (bits beyond the modulation's and bins without data point at the
     * first float after the symbol's LLRs): one 16-byte load per bin in demod.cu */
    for (int m = 0; m < W_NMAPS; m++) {
        const bool legacy = m < 4;
        const int nb = legacy ? nbl[m] : nbn[(m - 4) % 5];
        const int nsd = legacy ? 48 : 52;
        const int total = (m >= 9 ? 2 : 1) * nsd * nb;
        for (int i = 0; i < 64; i++) {
            const int j = legacy ? t->bin_j_l[i] : t->bin_j_nl[i];
            for (int b = 0; b < 8; b++) t->binmap[m][i][b] = (uint16_t)((j >= 0 && b < nb) ? t->map[m][j * nb + b] : total);
        }
    }
}
