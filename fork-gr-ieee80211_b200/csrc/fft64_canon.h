/* fft64_canon.h -- the 64-point forward DFT of this project, as ONE fixed
 * sequence of IEEE-754 binary32 operations.
 *
 * Why it exists.  The reference calls gr::fft::fft_complex_fwd(64,1) (FFTW,
 * version unpinned) at /root/reference/lib/demod_impl.cc:541-547 and
 * demod2_impl.cc:790-796.  FFTW is not available in this image, so both the
 * CPU oracle (oracle/mockgr/gnuradio/fft/fft.h) and the CUDA kernels
 * (demod.cuh / hdr.cuh) take their FFT from this header.  Every add, subtract
 * and multiply below is rounded on its own (no FMA contraction: the device
 * build maps the C8_* macros to __fadd_rn/__fsub_rn/__fmul_rn, the host build
 * must use -ffp-contract=off), so host and device produce bit-identical bins.
 *
 * Definition: X[k] = sum_n x[n] exp(-2*pi*i*n*k/64), unnormalised, natural
 * order in and out (bin 0 = DC), exactly what the reference expects.
 *
 * Algorithm: 64 = 8 x 8 Cooley-Tukey.  With n = 8*n1 + n2 and k = k1 + 8*k2:
 *   A[n2][k1] = sum_{n1} x[8*n1+n2] W8^(n1*k1)        (8 column FFTs)
 *   B[n2][k1] = A[n2][k1] * W64^(n2*k1)               (twiddle)
 *   X[k1+8*k2] = sum_{n2} B[n2][k1] W8^(n2*k2)        (8 row FFTs)
 * On the GPU one lane owns one n2 (then one k1) and c8_fft8 runs in registers.
 */
#ifndef W80211_FFT64_CANON_H
#define W80211_FFT64_CANON_H

#if defined(__CUDA_ARCH__)
#define C8_ADD(a, b) __fadd_rn((a), (b))
#define C8_SUB(a, b) __fsub_rn((a), (b))
#define C8_MUL(a, b) __fmul_rn((a), (b))
#else
#define C8_ADD(a, b) ((a) + (b))
#define C8_SUB(a, b) ((a) - (b))
#define C8_MUL(a, b) ((a) * (b))
#endif

#if defined(__CUDACC__)
#define C8_HD __host__ __device__ __forceinline__
#define C8_CONST static __device__ __constant__ const
#else
#define C8_HD static inline
#define C8_CONST static const
#endif

#define C8_SQRT1_2 0.707106769f /* (float)sqrt(0.5) */

/* In-place 8-point forward DFT, natural order in, natural order out. */
C8_HD void c8_fft8(float* re, float* im)
{
    /* radix-2 DIF stage: b[j] = a[j]+a[j+4]; b[j+4] = (a[j]-a[j+4]) * W8^j */
    float b0r = C8_ADD(re[0], re[4]), b0i = C8_ADD(im[0], im[4]);
    float b1r = C8_ADD(re[1], re[5]), b1i = C8_ADD(im[1], im[5]);
    float b2r = C8_ADD(re[2], re[6]), b2i = C8_ADD(im[2], im[6]);
    float b3r = C8_ADD(re[3], re[7]), b3i = C8_ADD(im[3], im[7]);
    float b4r = C8_SUB(re[0], re[4]), b4i = C8_SUB(im[0], im[4]);
    float t5r = C8_SUB(re[1], re[5]), t5i = C8_SUB(im[1], im[5]);
    float t6r = C8_SUB(re[2], re[6]), t6i = C8_SUB(im[2], im[6]);
    float t7r = C8_SUB(re[3], re[7]), t7i = C8_SUB(im[3], im[7]);
    /* W8^1 = (1-i)/sqrt2 : (x+iy) -> ((x+y) + i(y-x))/sqrt2 */
    float b5r = C8_MUL(C8_ADD(t5r, t5i), C8_SQRT1_2);
    float b5i = C8_MUL(C8_SUB(t5i, t5r), C8_SQRT1_2);
    /* W8^2 = -i : (x+iy) -> (y - ix) */
    float b6r = t6i, b6i = -t6r;
    /* W8^3 = (-1-i)/sqrt2 : (x+iy) -> ((y-x) - i(x+y))/sqrt2 */
    float b7r = C8_MUL(C8_SUB(t7i, t7r), C8_SQRT1_2);
    float b7i = -C8_MUL(C8_ADD(t7r, t7i), C8_SQRT1_2);

    /* 4-point DFT of (b0,b1,b2,b3) -> X0,X2,X4,X6 */
    float d0r = C8_ADD(b0r, b2r), d0i = C8_ADD(b0i, b2i);
    float d1r = C8_ADD(b1r, b3r), d1i = C8_ADD(b1i, b3i);
    float d2r = C8_SUB(b0r, b2r), d2i = C8_SUB(b0i, b2i);
    float e3r = C8_SUB(b1r, b3r), e3i = C8_SUB(b1i, b3i);
    float d3r = e3i, d3i = -e3r; /* times -i */
    re[0] = C8_ADD(d0r, d1r); im[0] = C8_ADD(d0i, d1i);
    re[4] = C8_SUB(d0r, d1r); im[4] = C8_SUB(d0i, d1i);
    re[2] = C8_ADD(d2r, d3r); im[2] = C8_ADD(d2i, d3i);
    re[6] = C8_SUB(d2r, d3r); im[6] = C8_SUB(d2i, d3i);

    /* 4-point DFT of (b4,b5,b6,b7) -> X1,X3,X5,X7 */
    float f0r = C8_ADD(b4r, b6r), f0i = C8_ADD(b4i, b6i);
    float f1r = C8_ADD(b5r, b7r), f1i = C8_ADD(b5i, b7i);
    float f2r = C8_SUB(b4r, b6r), f2i = C8_SUB(b4i, b6i);
    float g3r = C8_SUB(b5r, b7r), g3i = C8_SUB(b5i, b7i);
    float f3r = g3i, f3i = -g3r; /* times -i */
    re[1] = C8_ADD(f0r, f1r); im[1] = C8_ADD(f0i, f1i);
    re[5] = C8_SUB(f0r, f1r); im[5] = C8_SUB(f0i, f1i);
    re[3] = C8_ADD(f2r, f3r); im[3] = C8_ADD(f2i, f3i);
    re[7] = C8_SUB(f2r, f3r); im[7] = C8_SUB(f2i, f3i);
}

/* W64^m = exp(-2*pi*i*m/64), m = 0..63, each component rounded from double. */
C8_CONST float C8_W64_RE[64] = {
    1.0f, 0.99518472f, 0.980785251f, 0.956940353f,
    0.923879504f, 0.881921291f, 0.831469595f, 0.773010433f,
    0.707106769f, 0.634393275f, 0.555570245f, 0.471396744f,
    0.382683426f, 0.290284663f, 0.195090324f, 0.0980171412f,
    0.0f, -0.0980171412f, -0.195090324f, -0.290284663f,
    -0.382683426f, -0.471396744f, -0.555570245f, -0.634393275f,
    -0.707106769f, -0.773010433f, -0.831469595f, -0.881921291f,
    -0.923879504f, -0.956940353f, -0.980785251f, -0.99518472f,
    -1.0f, -0.99518472f, -0.980785251f, -0.956940353f,
    -0.923879504f, -0.881921291f, -0.831469595f, -0.773010433f,
    -0.707106769f, -0.634393275f, -0.555570245f, -0.471396744f,
    -0.382683426f, -0.290284663f, -0.195090324f, -0.0980171412f,
    0.0f, 0.0980171412f, 0.195090324f, 0.290284663f,
    0.382683426f, 0.471396744f, 0.555570245f, 0.634393275f,
    0.707106769f, 0.773010433f, 0.831469595f, 0.881921291f,
    0.923879504f, 0.956940353f, 0.980785251f, 0.99518472f,
};

C8_CONST float C8_W64_IM[64] = {
    0.0f, -0.0980171412f, -0.195090324f, -0.290284663f,
    -0.382683426f, -0.471396744f, -0.555570245f, -0.634393275f,
    -0.707106769f, -0.773010433f, -0.831469595f, -0.881921291f,
    -0.923879504f, -0.956940353f, -0.980785251f, -0.99518472f,
    -1.0f, -0.99518472f, -0.980785251f, -0.956940353f,
    -0.923879504f, -0.881921291f, -0.831469595f, -0.773010433f,
    -0.707106769f, -0.634393275f, -0.555570245f, -0.471396744f,
    -0.382683426f, -0.290284663f, -0.195090324f, -0.0980171412f,
    0.0f, 0.0980171412f, 0.195090324f, 0.290284663f,
    0.382683426f, 0.471396744f, 0.555570245f, 0.634393275f,
    0.707106769f, 0.773010433f, 0.831469595f, 0.881921291f,
    0.923879504f, 0.956940353f, 0.980785251f, 0.99518472f,
    1.0f, 0.99518472f, 0.980785251f, 0.956940353f,
    0.923879504f, 0.881921291f, 0.831469595f, 0.773010433f,
    0.707106769f, 0.634393275f, 0.555570245f, 0.471396744f,
    0.382683426f, 0.290284663f, 0.195090324f, 0.0980171412f,
};

/* B = A * W64^m with four separately rounded products. */
C8_HD void c8_twiddle(float* re, float* im, int m)
{
    float wr = C8_W64_RE[m & 63], wi = C8_W64_IM[m & 63];
    float ar = *re, ai = *im;
    *re = C8_SUB(C8_MUL(ar, wr), C8_MUL(ai, wi));
    *im = C8_ADD(C8_MUL(ar, wi), C8_MUL(ai, wr));
}

#if !defined(__CUDA_ARCH__)
/* Whole transform on one thread (host oracle; also usable in device debug code).
 * in/out: 64 interleaved (re,im) pairs; in and out may alias. */
static inline void c8_fft64(const float* in, float* out)
{
    float br[8][8], bi[8][8]; /* [n2][k1] */
    for (int n2 = 0; n2 < 8; n2++) {
        float r[8], i[8];
        for (int n1 = 0; n1 < 8; n1++) {
            r[n1] = in[2 * (8 * n1 + n2)];
            i[n1] = in[2 * (8 * n1 + n2) + 1];
        }
        c8_fft8(r, i);
        for (int k1 = 0; k1 < 8; k1++) {
            if (n2 * k1) c8_twiddle(&r[k1], &i[k1], n2 * k1);
            br[n2][k1] = r[k1];
            bi[n2][k1] = i[k1];
        }
    }
    for (int k1 = 0; k1 < 8; k1++) {
        float r[8], i[8];
        for (int n2 = 0; n2 < 8; n2++) {
            r[n2] = br[n2][k1];
            i[n2] = bi[n2][k1];
        }
        c8_fft8(r, i);
        for (int k2 = 0; k2 < 8; k2++) {
            out[2 * (k1 + 8 * k2)] = r[k2];
            out[2 * (k1 + 8 * k2) + 1] = i[k2];
        }
    }
}
#endif

#endif /* W80211_FFT64_CANON_H */
