/* demod_port.c -- plain-C restatement of the reference's per-symbol demodulation (one antenna), for checking only.
 *
 * TEST INFRASTRUCTURE.  Not part of the product: nothing under fork-gr-ieee80211_b200/, gr/ or include/ links, loads or
 * calls it.  Used by tests/test_port.py as a second opinion beside oracle/_ref, and as the CPU checker of the demod
 * half's data symbols on a machine where the reference tree is absent.
 *
 * What it restates (reference = /root/reference/lib/demod_impl.cc, cloud80211phy.cc):
 *   port_chan_nl          nonLegacyChanEstimate, one stream            demod_impl.cc:380-392
 *                         (HT/VHT-LTF signs LTF_NL_28_F_FLOAT          cloud80211phy.cc:113-129)
 *   port_demod_data       legacyChanUpdate / nonLegacyChanUpdate       demod_impl.cc:507-539, :413-447
 *                         fftDemod, pilotShift                         demod_impl.cc:541-557
 *                         pilot polarity PILOT_P, base patterns        cloud80211phy.cc:160-174
 *                         procSymQamToLlr                              cloud80211phy.cc:2090-2148
 *                         procSymDeintL2 / procSymDeintNL2SS1          cloud80211phy.cc:2150-2287
 * The header stage (format detection, SIG parsing) is NOT restated: the caller passes format, modulation and symbol
 * count, which the tests take from the packet generator or from the reference's output tags.
 * Tables are generated from the IEEE 802.11-2016 formulas (17.3.5.7, 17.3.5.10, 19.3.9.4.6, 19.3.11.8.3), not listed.
 * The FFT is fft64_canon.h, the same fixed sequence of binary32 operations the mock GNU Radio FFT of oracle/_ref uses
 * (FFTW is not in the reference tree).  Complex arithmetic is C99 float _Complex, i.e. libgcc's
 * __mulsc3 / __divsc3 and glibc's cabsf, exactly what std::complex<float> compiles to in the reference.
 * Pinned by tests/test_port.py against the reference's golden fixture (tests/golden/fixture_siso.npz: samples + channel
 * -> the LLR stream of the reference's demod block) and against oracle/_ref on generated legacy / HT / VHT packets.
 */
#include <complex.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "fft64_canon.h"

typedef float _Complex cf;

/* p_0..p_126 (17.3.5.10): the scrambler's output for the all-ones state, 0 -> +1, 1 -> -1 */
static void pilot_polarity(float* p)
{
    int st = 0x7f;
    for (int i = 0; i < 127; i++) {
        int fb = ((st >> 6) & 1) ^ ((st >> 3) & 1);
        st = ((st << 1) & 0x7e) | fb;
        p[i] = fb ? -1.0f : 1.0f;
    }
}

/* HT/VHT-LTF sign of FFT bin i (19.3.9.4.6: the L-LTF sequence with {1,1} added below and {-1,-1} above), 0 on unused bins */
static float ltf_nl_sign(int bin)
{
    static const char* lltf = "++--++-+-++++++--++-+-++++0+--++-+-+-----++--+-+-++++"; /* subcarriers -26..26, Equation 17-8 */
    int k = bin < 32 ? bin : bin - 64;
    if (k < -28 || k > 28 || k == 0) return 0.0f;
    if (k == -28 || k == -27) return 1.0f;
    if (k == 27 || k == 28) return -1.0f;
    return lltf[k + 26] == '+' ? 1.0f : -1.0f;
}

/* interleaver: coded bit k -> transmitted position (17.3.5.7 for legacy, 19.3.11.8.3 for HT/VHT 20 MHz, one stream) */
static int interleave_pos(int k, int ncbps, int nbpscs, int legacy)
{
    int s = nbpscs / 2 > 1 ? nbpscs / 2 : 1;
    int ncol = legacy ? 16 : 13;
    int nrow = legacy ? ncbps / 16 : 4 * nbpscs;
    int i = nrow * (k % ncol) + k / ncol;
    return s * (i / s) + (i + ncbps - (ncol * i) / ncbps) % s;
}

/* FFT of the 64 samples that start 8 samples into the symbol (C8P_SYM_SAMP_SHIFT, demod_impl.cc:541-547) */
static void fft_sym(const float* sym /* interleaved re,im */, cf* out)
{
    float o[128];
    c8_fft64(sym + 2 * 8, o);
    for (int i = 0; i < 64; i++) out[i] = CMPLXF(o[2 * i], o[2 * i + 1]);
}

/* nonLegacyChanEstimate, nSS = 1, nLTF = 1 (demod_impl.cc:380-392): H[i] = FFT(LTF)[i] / L_i on the 56 used bins */
void port_chan_nl(const float* ltf_sym, float* H_out /* 64 x (re, im) */)
{
    cf f[64];
    fft_sym(ltf_sym, f);
    memset(H_out, 0, 128 * sizeof(float));
    for (int i = 0; i < 64; i++) {
        if (i == 0 || (i >= 29 && i <= 35)) continue;
        cf h = f[i] / ltf_nl_sign(i);
        H_out[2 * i] = crealf(h);
        H_out[2 * i + 1] = cimagf(h);
    }
}

/* procSymQamToLlr (cloud80211phy.cc:2090-2148): I bits then Q bits of every subcarrier */
static void qam_to_llr(cf* q, float* llr, int nsd, int nbpscs)
{
    for (int i = 0; i < nsd; i++) {
        float* o = llr + i * nbpscs;
        switch (nbpscs) {
        case 1: o[0] = crealf(q[i]); break;
        case 2:
            q[i] *= 1.4142135623730951f;
            o[0] = crealf(q[i]);
            o[1] = cimagf(q[i]);
            break;
        case 4:
            q[i] *= 3.1622776601683795f;
            o[0] = crealf(q[i]);
            o[1] = 2.0f - fabsf(crealf(q[i]));
            o[2] = cimagf(q[i]);
            o[3] = 2.0f - fabsf(cimagf(q[i]));
            break;
        case 6:
            q[i] *= 6.48074069840786f;
            o[0] = crealf(q[i]);
            o[1] = 4.0f - fabsf(crealf(q[i]));
            o[2] = 2 - fabsf(4.0f - fabsf(crealf(q[i])));
            o[3] = cimagf(q[i]);
            o[4] = 4.0f - fabsf(cimagf(q[i]));
            o[5] = 2 - fabsf(4.0f - fabsf(cimagf(q[i])));
            break;
        default:
            q[i] *= 13.038404810405298f;
            o[0] = crealf(q[i]);
            o[1] = 8.0f - fabsf(crealf(q[i]));
            o[2] = 4 - fabsf(8.0f - fabsf(crealf(q[i])));
            o[3] = 2 - fabsf(4 - fabsf(8.0f - fabsf(crealf(q[i]))));
            o[4] = cimagf(q[i]);
            o[5] = 8.0f - fabsf(cimagf(q[i]));
            o[6] = 4 - fabsf(8.0f - fabsf(cimagf(q[i])));
            o[7] = 2 - fabsf(4 - fabsf(8.0f - fabsf(cimagf(q[i]))));
            break;
        }
    }
}

/* The data symbols of one packet on one antenna -> the LLR stream demod writes (coded-bit order, still punctured).
 * fmt: 0 legacy, 1 HT, 2 VHT.  sym0: first sample of the first data symbol.  nsymsamp: 80, or 72 with short GI.
 * H: legacy channel (tag "chan") for fmt 0, port_chan_nl's output otherwise.  Returns the number of LLRs, < 0 on bad arguments. */
int port_demod_data(int fmt, int nsym, int nsymsamp, int nbpscs, const float* sym0, const float* H_in, float* llr)
{
    if (fmt < 0 || fmt > 2 || nsym < 0 || (nsymsamp != 80 && nsymsamp != 72)) return -1;
    if (nbpscs != 1 && nbpscs != 2 && nbpscs != 4 && nbpscs != 6 && !(nbpscs == 8 && fmt == 2)) return -1;
    const int legacy = fmt == 0, nsd = legacy ? 48 : 52, ncbps = nsd * nbpscs;
    const int guard_lo = legacy ? 27 : 29, guard_hi = legacy ? 37 : 35;
    float pp[127];
    pilot_polarity(pp);
    /* demod_impl.cc:214,191,164: first polarity index 1 / 3 / 4; base pattern {1,1,1,-1} for all three formats */
    int pidx = legacy ? 1 : (fmt == 1 ? 3 : 4);
    float pilot[4] = {1.0f, 1.0f, 1.0f, -1.0f};
    int map[416];
    for (int k = 0; k < ncbps; k++) map[interleave_pos(k, ncbps, nbpscs, legacy)] = k; /* out[map[i]] = in[i] */
    cf H[64];
    for (int i = 0; i < 64; i++) H[i] = CMPLXF(H_in[2 * i], H_in[2 * i + 1]);
    for (int n = 0; n < nsym; n++) {
        cf f[64], e[64], q[52];
        float sl[416];
        fft_sym(sym0 + 2 * (size_t)n * nsymsamp, f);
        for (int i = 0; i < 64; i++) {
            if (i == 0 || (i >= guard_lo && i <= guard_hi)) continue;
            e[i] = f[i] / H[i];
        }
        /* bins 7, 21, 43 (-21), 57 (-7) against pattern entries 2, 3, 0, 1 (demod_impl.cc:426-430, :520-524) */
        cf sum = conjf(e[7] * pilot[2] * pp[pidx] + e[21] * pilot[3] * pp[pidx] + e[43] * pilot[0] * pp[pidx] + e[57] * pilot[1] * pp[pidx]);
        if (!legacy) { /* pilotShift, demod_impl.cc:549-557 */
            float t = pilot[0];
            pilot[0] = pilot[1];
            pilot[1] = pilot[2];
            pilot[2] = pilot[3];
            pilot[3] = t;
        }
        pidx = (pidx + 1) % 127;
        const float a = cabsf(sum);
        int j = nsd / 2; /* ascending frequency: bin 1 is subcarrier nsd/2 */
        for (int i = 0; i < 64; i++) {
            if (i == 0 || (i >= guard_lo && i <= guard_hi) || i == 7 || i == 21 || i == 43 || i == 57) continue;
            q[j] = e[i] * sum / a;
            if (++j >= nsd) j = 0;
        }
        qam_to_llr(q, sl, nsd, nbpscs);
        float* out = llr + (size_t)n * ncbps;
        for (int i = 0; i < ncbps; i++) out[map[i]] = sl[i];
    }
    return nsym * ncbps;
}
