// divcheck.cu -- brute-force check of the known-divisor division of csrc/common.cuh against the plain IEEE
// division on the GPU: d_div_known vs __ddiv_rn (binary64) and f_div_known vs __fdiv_rn (binary32).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -I../fork-gr-ieee80211_b200/csrc -o divcheck divcheck.cu
// Prints the number of operand pairs tested and the number of mismatches (must be 0).
#include <cstdio>
#include "common.cuh"

__device__ __forceinline__ unsigned long long rng(unsigned long long& s)
{
    s ^= s << 13; s ^= s >> 7; s ^= s << 17;
    return s;
}
__device__ __forceinline__ float rnd_float(unsigned long long& s, int emin, int emax)
{
    unsigned long long r = rng(s);
    unsigned int mant = (unsigned int)(r & 0x7fffffu);
    int e = emin + (int)((r >> 23) % (unsigned)(emax - emin + 1));
    unsigned int sign = (unsigned int)((r >> 60) & 1u) << 31;
    return __uint_as_float(sign | ((unsigned)(e + 127) << 23) | mant);
}

__global__ void check(unsigned long long seed, int iters, unsigned long long* bad, unsigned long long* n)
{
    unsigned long long s = seed + 0x9E3779B97F4A7C15ull * (blockIdx.x * blockDim.x + threadIdx.x + 1);
    unsigned long long nb = 0, nn = 0;
    for (int it = 0; it < iters; it++) {
        /* the demod's use: a, b (received bin), c, d (channel tap) floats; denom = c*c + d*d */
        float a = rnd_float(s, -12, 6), b = rnd_float(s, -12, 6), c = rnd_float(s, -8, 4), d = rnd_float(s, -8, 4);
        if (it & 1) { /* hard cases: significands with long runs of ones / zeros */
            c = __uint_as_float(__float_as_uint(c) | 0x007ff000u);
            d = __uint_as_float(__float_as_uint(d) & 0xff800fffu);
        }
        cf32 num = c_make(a, b), den = c_make(c, d);
        double dc = c, dd = d;
        double denom = __dadd_rn(__dmul_rn(dc, dc), __dmul_rn(dd, dd));
        if (!d_recip_unsafe(denom)) {
            double y = __drcp_rn(denom);
            cf32 r1 = c_div(num, den), r2 = c_div_known(num, den, denom, y);
            nn += 2;
            nb += (__float_as_uint(r1.x) != __float_as_uint(r2.x)) + (__float_as_uint(r1.y) != __float_as_uint(r2.y));
            /* and the raw binary64 quotient */
            double na = __fma_rn((double)b, dd, __dmul_rn((double)a, dc));
            double q1 = __ddiv_rn(na, denom), q2 = d_div_known(na, denom, y);
            nn += 1;
            nb += (__double_as_longlong(q1) != __double_as_longlong(q2));
        }
        /* binary32: x / pa */
        float x = rnd_float(s, -20, 8), pa = fabsf(rnd_float(s, -6, 6));
        if (it & 2) pa = __uint_as_float(__float_as_uint(pa) | 0x007fff00u);
        if (!f_recip_unsafe(pa)) {
            float y = __frcp_rn(pa);
            float q1 = __fdiv_rn(x, pa), q2 = f_div_known(x, pa, y);
            nn += 1;
            nb += (__float_as_uint(q1) != __float_as_uint(q2));
        }
    }
    atomicAdd(bad, nb);
    atomicAdd(n, nn);
}

int main()
{
    unsigned long long *d, h[2] = {0, 0};
    cudaMalloc(&d, 16);
    cudaMemcpy(d, h, 16, cudaMemcpyHostToDevice);
    for (int r = 0; r < 8; r++) check<<<148 * 8, 256>>>(0x1234567ull + 7919ull * r, 4096, d, d + 1);
    cudaDeviceSynchronize();
    cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("divcheck: %llu quotients compared, %llu mismatches (%s)\n", h[1], h[0], cudaGetErrorString(cudaGetLastError()));
    return h[0] != 0;
}
