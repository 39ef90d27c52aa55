// Micro-benchmark: issue throughput of the integer instructions the SW kernel is made of, on sm_100a.
// Prints lane-ops per clock per SM for each mix, so the integer roofline in DESIGN.md is measured, not assumed.
#include <cuda_runtime.h>
#include <cstdio>

#define ITER 4096
template <int MODE> __global__ void k(unsigned *out, unsigned a0, unsigned b0)
{
    unsigned x0 = threadIdx.x + a0, x1 = x0 ^ 0x1111, x2 = x0 + 77, x3 = x0 * 3, x4 = x0 + 5, x5 = x0 ^ 9, x6 = x0 + 1, x7 = x0 - 1;
    const unsigned m1 = 0xffffffffu, b = b0;
#pragma unroll 1
    for (int i = 0; i < ITER; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (MODE == 0) {  // VIADDMNMX.S16x2.RELU
                x0 = __viaddmax_s16x2_relu(x0, m1, b); x1 = __viaddmax_s16x2_relu(x1, m1, b); x2 = __viaddmax_s16x2_relu(x2, m1, b); x3 = __viaddmax_s16x2_relu(x3, m1, b);
                x4 = __viaddmax_s16x2_relu(x4, m1, b); x5 = __viaddmax_s16x2_relu(x5, m1, b); x6 = __viaddmax_s16x2_relu(x6, m1, b); x7 = __viaddmax_s16x2_relu(x7, m1, b);
            } else if (MODE == 1) {  // VIMNMX3.S16x2
                x0 = __vimax3_s16x2(x0, x1, b); x1 = __vimax3_s16x2(x1, x2, b); x2 = __vimax3_s16x2(x2, x3, b); x3 = __vimax3_s16x2(x3, x4, b);
                x4 = __vimax3_s16x2(x4, x5, b); x5 = __vimax3_s16x2(x5, x6, b); x6 = __vimax3_s16x2(x6, x7, b); x7 = __vimax3_s16x2(x7, x0, b);
            } else if (MODE == 2) {  // PRMT
                x0 = __byte_perm(x0, b, x1); x1 = __byte_perm(x1, b, x2); x2 = __byte_perm(x2, b, x3); x3 = __byte_perm(x3, b, x4);
                x4 = __byte_perm(x4, b, x5); x5 = __byte_perm(x5, b, x6); x6 = __byte_perm(x6, b, x7); x7 = __byte_perm(x7, b, x0);
            } else if (MODE == 3) {  // IADD (whatever pipe ptxas picks)
                x0 += x1; x1 += x2; x2 += x3; x3 += x4; x4 += x5; x5 += x6; x6 += x7; x7 += x0;
            } else if (MODE == 4) {  // VIMNMX.S16x2 (2-input)
                x0 = __vmaxs2(x0, x1); x1 = __vmaxs2(x1, x2); x2 = __vmaxs2(x2, x3); x3 = __vmaxs2(x3, x4);
                x4 = __vmaxs2(x4, x5); x5 = __vmaxs2(x5, x6); x6 = __vmaxs2(x6, x7); x7 = __vmaxs2(x7, x0);
            } else if (MODE == 5) {  // the SW cell mix: PRMT + IADD + VIMNMX3 + VIADDMNMX (+ half a VIMNMX3), two independent chains
                unsigned s = __byte_perm(x2, b, x3); unsigned t = x0 + s; unsigned m = __vimax3_s16x2(t, x1, x4); x4 = __viaddmax_s16x2_relu(m, m1, 0u); x0 = x1; x1 = x4;
                unsigned s2 = __byte_perm(x3, b, x2); unsigned t2 = x5 + s2; unsigned m2 = __vimax3_s16x2(t2, x6, x7); x7 = __viaddmax_s16x2_relu(m2, m1, 0u); x5 = x6; x6 = x7;
                x2 = __vimax3_s16x2(x2, m, m2);
            } else if (MODE == 6) {  // VIADD.16x2
                x0 = __vadd2(x0, x1); x1 = __vadd2(x1, x2); x2 = __vadd2(x2, x3); x3 = __vadd2(x3, x4);
                x4 = __vadd2(x4, x5); x5 = __vadd2(x5, x6); x6 = __vadd2(x6, x7); x7 = __vadd2(x7, x0);
            } else if (MODE == 7) {  // IMAD (fma pipe) interleaved with VIMNMX3 (alu pipe)
                x0 = x0 * b + x1; x1 = __vimax3_s16x2(x1, x2, b); x2 = x2 * b + x3; x3 = __vimax3_s16x2(x3, x4, b);
                x4 = x4 * b + x5; x5 = __vimax3_s16x2(x5, x6, b); x6 = x6 * b + x7; x7 = __vimax3_s16x2(x7, x0, b);
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 ^ x1 ^ x2 ^ x3 ^ x4 ^ x5 ^ x6 ^ x7;
}

template <int MODE> void run(const char *name, double ops_per_inner, unsigned *out, int sms, double ghz)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int ctas = sms * 4, threads = 512;
    k<MODE><<<ctas, threads>>>(out, 1, 3);
    cudaEventRecord(a);
    k<MODE><<<ctas, threads>>>(out, 1, 3);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    double ops = (double)ctas * threads * ITER * 4 * ops_per_inner;
    printf("%-44s %8.3f ms  %8.2f Tlane-op/s  %6.1f lane-ops/clk/SM (at %.3f GHz)\n", name, ms, ops / ms / 1e9, ops / (ms * 1e-3) / sms / (ghz * 1e9), ghz);
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double ghz = khz / 1e6;
    printf("%s, %d SMs, max clock %.3f GHz\n", p.name, p.multiProcessorCount, ghz);
    unsigned *out;
    cudaMalloc(&out, 148 * 8 * 512 * 4);
    int sms = p.multiProcessorCount;
    run<0>("VIADDMNMX.S16x2.RELU", 8, out, sms, ghz);
    run<1>("VIMNMX3.S16x2", 8, out, sms, ghz);
    run<2>("PRMT", 8, out, sms, ghz);
    run<3>("IADD", 8, out, sms, ghz);
    run<4>("VIMNMX.S16x2", 8, out, sms, ghz);
    run<6>("VIADD.16x2", 8, out, sms, ghz);
    run<7>("IMAD + VIMNMX3 interleaved", 8, out, sms, ghz);
    run<5>("SW cell mix (9 instr = 2 cell pairs)", 9, out, sms, ghz);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
