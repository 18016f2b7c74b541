// pipebench.cu -- issue-rate microbenchmark for the instruction mix of the Viterbi ACS on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipebench pipebench.cu ; run on the GPU box.
// Prints warp-instructions per clock per SM for each op (4 SMSPs: 4.0 = one per clock per scheduler).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
#define ITER 2048
#define NCH 8

template <int OP>
__global__ void __launch_bounds__(512) k(float* out, const float* in, u64* cyc, int n_iter)
{
    float a[NCH], b[NCH];
    u64 A[NCH], B[NCH];
    unsigned ua[NCH], ub[NCH];
    for (int i = 0; i < NCH; i++) {
        a[i] = in[threadIdx.x + i]; b[i] = in[threadIdx.x + 32 + i];
        ua[i] = __float_as_uint(a[i]); ub[i] = __float_as_uint(b[i]);
        asm("mov.b64 %0, {%1,%2};" : "=l"(A[i]) : "f"(a[i]), "f"(b[i]));
        asm("mov.b64 %0, {%1,%2};" : "=l"(B[i]) : "f"(b[i]), "f"(a[i]));
    }
    __syncthreads();
    u64 t0 = clock64();
    for (int it = 0; it < n_iter; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
#pragma unroll
            for (int i = 0; i < NCH; i++) {
                if (OP == 0) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(b[i]));
                if (OP == 1) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(A[i]) : "l"(B[i]));
                if (OP == 2) asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(b[i]));
                if (OP == 3) asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(ua[i]) : "r"(ub[i]));
                if (OP == 4) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(ua[i]) : "r"(ub[i]), "r"(ua[(i + 1) % NCH]));
                if (OP == 5) asm volatile("mad.lo.u32 %0, %0, %1, %1;" : "+r"(ua[i]) : "r"(ub[i]));
                if (OP == 6) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(ua[i]) : "r"(ub[i]));
                if (OP == 7) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(A[i]) : "l"(B[i]));
                if (OP == 8) asm volatile("prmt.b32 %0, %0, %1, 0x7531;" : "+r"(ua[i]) : "r"(ub[i]));
                if (OP == 9) { /* packed butterfly pair: 6 FADD2 + 4 FMNMX + 4 SHF */
                    u64 c0, c1, e0, e1, dl, dh;
                    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(c0) : "l"(A[i]), "l"(B[0]));
                    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(c1) : "l"(A[(i + 1) % NCH]), "l"(B[1]));
                    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(e0) : "l"(A[i]), "l"(B[1]));
                    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(e1) : "l"(A[(i + 1) % NCH]), "l"(B[0]));
                    asm volatile("sub.rn.f32x2 %0, %1, %2;" : "=l"(dl) : "l"(c0), "l"(c1));
                    asm volatile("sub.rn.f32x2 %0, %1, %2;" : "=l"(dh) : "l"(e0), "l"(e1));
                    float x0, x1, y0, y1, z0, z1, w0, w1, p0, p1, q0, q1, m0, m1, m2, m3;
                    asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(x0), "=f"(x1) : "l"(c0));
                    asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(y0), "=f"(y1) : "l"(c1));
                    asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(z0), "=f"(z1) : "l"(e0));
                    asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(w0), "=f"(w1) : "l"(e1));
                    asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(p0), "=f"(p1) : "l"(dl));
                    asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(q0), "=f"(q1) : "l"(dh));
                    asm volatile("max.f32 %0, %1, %2;" : "=f"(m0) : "f"(x0), "f"(y0));
                    asm volatile("max.f32 %0, %1, %2;" : "=f"(m1) : "f"(x1), "f"(y1));
                    asm volatile("max.f32 %0, %1, %2;" : "=f"(m2) : "f"(z0), "f"(w0));
                    asm volatile("max.f32 %0, %1, %2;" : "=f"(m3) : "f"(z1), "f"(w1));
                    asm volatile("mov.b64 %0, {%1,%2};" : "=l"(A[i]) : "f"(m0), "f"(m1));
                    asm volatile("mov.b64 %0, {%1,%2};" : "=l"(A[(i + 1) % NCH]) : "f"(m2), "f"(m3));
                    asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(ua[0]) : "r"(__float_as_uint(p1)));
                    asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(ua[0]) : "r"(__float_as_uint(p0)));
                    asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(ua[1]) : "r"(__float_as_uint(q1)));
                    asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(ua[1]) : "r"(__float_as_uint(q0)));
                }
                if (OP == 10) { /* scalar butterfly: 6 FADD + 2 FMNMX + 2 SHF */
                    float c0, c1, e0, e1, dl, dh;
                    asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(c0) : "f"(a[i]), "f"(b[0]));
                    asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(c1) : "f"(a[(i + 1) % NCH]), "f"(b[1]));
                    asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(e0) : "f"(a[i]), "f"(b[1]));
                    asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(e1) : "f"(a[(i + 1) % NCH]), "f"(b[0]));
                    asm volatile("sub.rn.f32 %0, %1, %2;" : "=f"(dl) : "f"(c0), "f"(c1));
                    asm volatile("sub.rn.f32 %0, %1, %2;" : "=f"(dh) : "f"(e0), "f"(e1));
                    asm volatile("max.f32 %0, %1, %2;" : "=f"(a[i]) : "f"(c0), "f"(c1));
                    asm volatile("max.f32 %0, %1, %2;" : "=f"(a[(i + 1) % NCH]) : "f"(e0), "f"(e1));
                    asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(ua[0]) : "r"(__float_as_uint(dl)));
                    asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(ua[1]) : "r"(__float_as_uint(dh)));
                }
                if (OP == 11) { /* FMNMX + FADD2 alternating 2:3 (ALU:FMA ratio of the packed butterfly without SHF) */
                    asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(b[i]));
                    asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(A[i]) : "l"(B[i]));
                }
                if (OP == 12) { /* setp + selp */
                    asm volatile("{.reg .pred p; setp.gt.f32 p, %0, %1; selp.f32 %0, %0, %1, p;}" : "+f"(a[i]) : "f"(b[i]));
                }
                if (OP == 13) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(A[i]) : "l"(B[i]));
                if (OP == 14) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(a[i]) : "f"(b[i]));
                if (OP == 15) { /* FMNMX + SHF pair (both ALU?) */
                    asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(b[i]));
                    asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(ua[i]) : "r"(ub[i]));
                }
                if (OP == 16) { /* FADD + SHF pair (FMA + ALU) */
                    asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(b[i]));
                    asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(ua[i]) : "r"(ub[i]));
                }
                if (OP == 17) asm volatile("min.s32 %0, %0, %1;" : "+r"(ua[i]) : "r"(ub[i]));
                if (OP == 18) asm volatile("add.s32 %0, %0, %1;" : "+r"(ua[i]) : "r"(ub[i]));
                if (OP == 19) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(b[i]), "f"(a[(i + 1) % NCH]));
            }
        }
    }
    u64 t1 = clock64();
    float s = 0; unsigned us = 0;
    for (int i = 0; i < NCH; i++) { float x, y; asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(A[i])); s += a[i] + x + y; us ^= ua[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + __uint_as_float(us);
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int per_inner, float* out, float* in, u64* cyc, int nsm)
{
    const int n_iter = ITER;
    k<OP><<<nsm, 512>>>(out, in, cyc, 16);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<OP><<<nsm, 512>>>(out, in, cyc, n_iter);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    u64 h[256]; cudaMemcpy(h, cyc, nsm * sizeof(u64), cudaMemcpyDeviceToHost);
    double c = 0; for (int i = 0; i < nsm; i++) c += (double)h[i]; c /= nsm;
    double winst = (double)n_iter * 8 * NCH * per_inner * 16; /* 16 warps per SM */
    printf("%-34s %7.3f warp-inst/clk/SM   (%.0f cycles, %.3f ms, SM clock %.0f MHz)%s\n", name, winst / c, c, ms, c / ms / 1e3,
           cudaGetLastError() == cudaSuccess ? "" : " ERROR");
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int nsm = p.multiProcessorCount;
    printf("%s, %d SMs\n", p.name, nsm);
    float *out, *in; u64* cyc;
    cudaMalloc(&out, nsm * 512 * 4); cudaMalloc(&in, 4096 * 4); cudaMalloc(&cyc, 256 * 8);
    float hin[4096]; for (int i = 0; i < 4096; i++) hin[i] = 1.0f + 1e-3f * i;
    cudaMemcpy(in, hin, sizeof(hin), cudaMemcpyHostToDevice);
    run<0>("FADD", 1, out, in, cyc, nsm);
    run<1>("FADD2", 1, out, in, cyc, nsm);
    run<14>("FFMA", 1, out, in, cyc, nsm);
    run<13>("FFMA2", 1, out, in, cyc, nsm);
    run<7>("FMUL2", 1, out, in, cyc, nsm);
    run<2>("FMNMX", 1, out, in, cyc, nsm);
    run<19>("FMNMX3", 1, out, in, cyc, nsm);
    run<3>("SHF.L.W", 1, out, in, cyc, nsm);
    run<4>("LOP3", 1, out, in, cyc, nsm);
    run<8>("PRMT", 1, out, in, cyc, nsm);
    run<17>("IMNMX", 1, out, in, cyc, nsm);
    run<18>("IADD", 1, out, in, cyc, nsm);
    run<5>("IMAD", 1, out, in, cyc, nsm);
    run<6>("IMAD.HI", 1, out, in, cyc, nsm);
    run<12>("FSETP+FSEL", 2, out, in, cyc, nsm);
    run<15>("FMNMX+SHF", 2, out, in, cyc, nsm);
    run<16>("FADD+SHF", 2, out, in, cyc, nsm);
    run<11>("FMNMX+FADD2", 2, out, in, cyc, nsm);
    run<9>("packed butterfly pair (14 inst)", 14, out, in, cyc, nsm);
    run<10>("scalar butterfly (10 inst)", 10, out, in, cyc, nsm);
    return 0;
}
