/* presiso_port.c -- CPU restatement of examples/presiso.grc of the reference (TEST INFRASTRUCTURE ONLY: only tests/ may call it).
 *
 * presiso is not reference source: it is a GRC hierarchy (/root/reference/examples/presiso.grc:30-230) of stock GNU Radio
 * blocks, a third-party dependency that is absent from /root/reference (GNU Radio 3.10, gr-blocks; version unpinned by the
 * reference, README asks for "GNU Radio 3.10"):
 *
 *   blocks_delay(16)                              presiso.grc:69-86
 *   blocks_multiply_conjugate_cc                  presiso.grc:146-160   out = in0 * conj(in1), in0 = delayed, in1 = direct
 *   blocks_moving_average_cc(48, 1, 4000)         presiso.grc:106-124   -> output 1 (preconj), complex_to_mag -> divide in0
 *   blocks_complex_to_mag_squared
 *   blocks_moving_average_ff(64, 1, 4000)         presiso.grc:126-144   -> divide in1
 *   blocks_divide_ff                              presiso.grc:88-104    -> output 0 (preac)
 *
 * Published algorithm of gr::blocks::moving_average (gr-blocks/lib/moving_average_impl.cc, vlen == 1), restated:
 *     history = length;  work(): n = min(noutput_items, max_iter);
 *         sum carried over:   for i < n:  sum += in[i + length - 1];  out[i] = sum * scale;  sum -= in[i];        (3.9 / 3.10)
 *         sum re-made:        sum = in[0] + ... + in[length - 2] first, then the same loop                        (3.7 / 3.8)
 * Both are sequential binary32 (complex: two binary32) running sums; the second depends on how the scheduler cuts the
 * stream into work() calls (at most max_iter = 4000 outputs each).  `chunk` selects: 0 = carried over the whole stream,
 * > 0 = re-made every `chunk` outputs.  The products of multiply_conjugate are VOLK kernels whose rounding depends on the
 * machine's SIMD level (with or without FMA); here: four rounded products, rounded sums (the generic kernel).
 * So presiso has no bit-exact definition across machines and schedulers; this file pins what CAN be pinned: the GPU
 * stage (csrc/detect.cu presiso_kernel) stays within 1e-4 of both variants, and the trigger / sync decisions taken on its
 * streams equal those taken on these (tests/test_gpu_detect.py).  Samples before the stream count as zero (the history
 * GNU Radio prepends). */
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* x: n complex samples (re, im); ac: n floats; cj: n complex */
void port_presiso(const float* x, long n, int chunk, float* ac, float* cj)
{
    /* product stream p[i] = x[i-16] * conj(x[i]) and power stream q[i] = |x[i]|^2, with 47 / 63 zeros of history in front */
    float* pr = (float*)calloc((size_t)(n + 64) * 2, sizeof(float));
    float* pw = (float*)calloc((size_t)(n + 64), sizeof(float));
    const int HP = 47, HQ = 63;
    for (long i = 0; i < n; i++) {
        const float br = x[2 * i], bi = x[2 * i + 1];
        float ar = 0.f, ai = 0.f;
        if (i >= 16) {
            ar = x[2 * (i - 16)];
            ai = x[2 * (i - 16) + 1];
        }
        /* (ar + j ai)(br - j bi) */
        const float rr = ar * br, ii = ai * bi, ri = ai * br, ir = ar * bi;
        pr[2 * (i + HP)] = rr + ii;
        pr[2 * (i + HP) + 1] = ri - ir;
        pw[i + HQ] = br * br + bi * bi;
    }
    float sr = 0.f, si = 0.f, sp = 0.f;
    for (long i = 0; i < n; i++) {
        if (chunk > 0 && i % chunk == 0) {
            /* a new work() call: the sums are made again from the length - 1 items of history */
            sr = si = 0.f;
            for (int k = 0; k < 47; k++) {
                sr += pr[2 * (i + k)];
                si += pr[2 * (i + k) + 1];
            }
            sp = 0.f;
            for (int k = 0; k < 63; k++) sp += pw[i + k];
        }
        sr += pr[2 * (i + 47)];
        si += pr[2 * (i + 47) + 1];
        sp += pw[i + 63];
        cj[2 * i] = sr;
        cj[2 * i + 1] = si;
        /* complex_to_mag: sqrt(re^2 + im^2) in binary32 (volk_32fc_magnitude_32f generic) */
        const float mag = sqrtf(sr * sr + si * si);
        ac[i] = mag / sp;
        sr -= pr[2 * i];
        si -= pr[2 * i + 1];
        sp -= pw[i];
    }
    free(pr);
    free(pw);
}
