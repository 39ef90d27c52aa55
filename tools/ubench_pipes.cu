// Micro-benchmark #3: register-operand effects and pipe assignment of the ops in the SW recurrence (sm_100a).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_pipes tools/ubench_pipes.cu
// Output: warp instructions per clock per SM (4.0 = issue limit) for each loop body.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>

#define ITER 4096

__device__ __forceinline__ unsigned hmax2u(unsigned a, unsigned b)
{
    unsigned r;
    asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ unsigned hfma_relu(unsigned a, unsigned b, unsigned c)
{
    unsigned r;
    asm("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ unsigned hadd2u(unsigned a, unsigned b)
{
    unsigned r;
    asm("add.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

template <int MODE> __global__ void __launch_bounds__(256) k(unsigned *out, unsigned a0, unsigned b0)
{
    unsigned x[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) x[i] = (threadIdx.x * (2 * i + 1) + a0 + i) & 0x00ff00ffu;
    const unsigned b = b0, one = 0x3c003c00u, negulp = 0x80018001u;
#pragma unroll 1
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (MODE == 0) {  // VIMNMX3, two register operands + constant
#pragma unroll
                for (int i = 0; i < 8; ++i) x[i] = __vimax3_s16x2(x[i], x[(i + 1) & 7], b);
            } else if (MODE == 1) {  // VIMNMX3, three distinct register operands
#pragma unroll
                for (int i = 0; i < 8; ++i) x[i] = __vimax3_s16x2(x[i], x[(i + 3) & 7], x[(i + 5) & 7]);
            } else if (MODE == 2) {  // VIMNMX (2-input), two register operands
#pragma unroll
                for (int i = 0; i < 8; ++i) x[i] = __vmaxs2(x[i], x[(i + 3) & 7]);
            } else if (MODE == 3) {  // HMNMX2 alone
#pragma unroll
                for (int i = 0; i < 8; ++i) x[i] = hmax2u(x[i], x[(i + 3) & 7]);
            } else if (MODE == 4) {  // HMNMX2 + VIMNMX (2-reg) interleaved: > 2.0/clk/SM means different pipes
#pragma unroll
                for (int i = 0; i < 8; i += 2) { x[i] = hmax2u(x[i], x[(i + 3) & 7]); x[i + 1] = __vmaxs2(x[i + 1], x[(i + 4) & 7]); }
            } else if (MODE == 5) {  // HMNMX2 + HFMA2 interleaved
#pragma unroll
                for (int i = 0; i < 8; i += 2) { x[i] = hmax2u(x[i], x[(i + 3) & 7]); x[i + 1] = hfma_relu(x[i + 1], one, x[(i + 4) & 7]); }
            } else if (MODE == 6) {  // HFMA2.RELU + VIMNMX (2-reg) interleaved: ideal 4.0
#pragma unroll
                for (int i = 0; i < 8; i += 2) { x[i] = hfma_relu(x[i], one, negulp); x[i + 1] = __vmaxs2(x[i + 1], x[(i + 4) & 7]); }
            } else if (MODE == 7) {  // HFMA2.RELU + VIMNMX3 (3-reg) interleaved
#pragma unroll
                for (int i = 0; i < 8; i += 2) { x[i] = hfma_relu(x[i], one, negulp); x[i + 1] = __vimax3_s16x2(x[i + 1], x[(i + 4) & 7], x[(i + 6) & 7]); }
            } else if (MODE == 8) {  // HADD2 + HFMA2.RELU + VIMNMX3(3-reg): the all-fma-clamp cell without lookups, 2 chains
                // chain A: x0 diag, x1 up, x2 left ; chain B: x4, x5, x6
                unsigned t = hadd2u(x[3], x[0]); unsigned m = __vimax3_s16x2(t, x[1], x[2]); x[2] = hfma_relu(m, one, negulp); x[0] = x[1]; x[1] = x[2];
                unsigned t2 = hadd2u(x[7], x[4]); unsigned m2 = __vimax3_s16x2(t2, x[5], x[6]); x[6] = hfma_relu(m2, one, negulp); x[4] = x[5]; x[5] = x[6];
                x[8] = __vimax3_s16x2(x[8], m, m2);
            } else if (MODE == 9) {  // same with 2-input maxes only: hmax2(hmax2(x, up), left) on whatever pipe HMNMX2 uses
                unsigned t = hadd2u(x[3], x[0]); unsigned m = hmax2u(hmax2u(t, x[1]), x[2]); x[2] = hfma_relu(m, one, negulp); x[0] = x[1]; x[1] = x[2];
                unsigned t2 = hadd2u(x[7], x[4]); unsigned m2 = hmax2u(hmax2u(t2, x[5]), x[6]); x[6] = hfma_relu(m2, one, negulp); x[4] = x[5]; x[5] = x[6];
                x[8] = __vimax3_s16x2(x[8], m, m2);
            } else if (MODE == 10) {  // 4 independent chains of the all-fma-clamp cell
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    unsigned *y = x + 3 * c;
                    unsigned t = hadd2u(b, y[0]); unsigned m = __vimax3_s16x2(t, y[1], y[2]); y[2] = hfma_relu(m, one, negulp); y[0] = y[1]; y[1] = y[2];
                }
            }
        }
    }
    unsigned r = 0;
#pragma unroll
    for (int i = 0; i < 12; ++i) r ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE> void run(const char *name, double instr_per_inner, int warps_per_sm, unsigned *out, int sms, double ghz)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int threads = 256, ctas = sms * (warps_per_sm * 32 / threads);
    k<MODE><<<ctas, threads>>>(out, 1, 3);
    cudaEventRecord(a);
    k<MODE><<<ctas, threads>>>(out, 1, 3);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    const double warps = (double)ctas * threads / 32, clk = ms * 1e-3 * ghz * 1e9;
    printf("%-78s %2d warps/SM %8.3f ms  %5.2f warp-instr/clk/SM\n", name, warps_per_sm, ms, warps * ITER * 4 * instr_per_inner / clk / sms);
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ghz = khz / 1e6;
    const int sms = p.multiProcessorCount;
    printf("%s, %d SMs, max clock %.3f GHz (rates assume the max clock)\n", p.name, sms, ghz);
    unsigned *out;
    cudaMalloc(&out, (size_t)sms * 64 * 32 * 4);
    for (int w : {16, 32}) {
        run<0>("VIMNMX3.S16x2 (2 regs + const)", 8, w, out, sms, ghz);
        run<1>("VIMNMX3.S16x2 (3 regs)", 8, w, out, sms, ghz);
        run<2>("VIMNMX.S16x2 (2 regs)", 8, w, out, sms, ghz);
        run<3>("HMNMX2", 8, w, out, sms, ghz);
        run<4>("HMNMX2 + VIMNMX 1:1", 8, w, out, sms, ghz);
        run<5>("HMNMX2 + HFMA2.RELU 1:1", 8, w, out, sms, ghz);
        run<6>("HFMA2.RELU + VIMNMX(2 regs) 1:1", 8, w, out, sms, ghz);
        run<7>("HFMA2.RELU + VIMNMX3(3 regs) 1:1", 8, w, out, sms, ghz);
        run<8>("cell: HADD2 + VIMNMX3(3r) + HFMA2.RELU, 2 chains (+best)  [7 instr = 2 cells]", 7, w, out, sms, ghz);
        run<9>("cell: HADD2 + 2 HMNMX2 + HFMA2.RELU, 2 chains (+best)      [9 instr = 2 cells]", 9, w, out, sms, ghz);
        run<10>("cell: HADD2 + VIMNMX3(3r) + HFMA2.RELU, 4 chains           [12 instr = 4 cells]", 12, w, out, sms, ghz);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
