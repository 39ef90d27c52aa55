// Micro-benchmark #4: which sub-pipes do IMAD / HFMA2 / HADD2 / FFMA / VIMNMX use on sm_100a, and how close to the
// 4 warp-instr/clk/SM issue limit do 2- and 3-pipe mixes of the SW recurrence's instructions get?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_pipes2 tools/ubench_pipes2.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>

#define ITER 4096

__device__ __forceinline__ unsigned hfma_relu(unsigned a, unsigned b, unsigned c)
{
    unsigned r;
    asm volatile("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ unsigned hadd2u(unsigned a, unsigned b)
{
    unsigned r;
    asm volatile("add.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ unsigned imad(unsigned a, unsigned b, unsigned c)
{
    unsigned r;
    asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ unsigned ffma(unsigned a, unsigned b, unsigned c)
{
    float r;
    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b)), "f"(__uint_as_float(c)));
    return __float_as_uint(r);
}
__device__ __forceinline__ unsigned fadd(unsigned a, unsigned b)
{
    float r;
    asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b)));
    return __float_as_uint(r);
}
__device__ __forceinline__ unsigned lop3(unsigned a, unsigned b, unsigned c)
{
    unsigned r;
    asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
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
            if (MODE == 0) {  // IMAD alone
#pragma unroll
                for (int i = 0; i < 8; ++i) x[i] = imad(x[i], b, x[(i + 3) & 7]);
            } else if (MODE == 1) {  // HFMA2 alone
#pragma unroll
                for (int i = 0; i < 8; ++i) x[i] = hfma_relu(x[i], one, x[(i + 3) & 7]);
            } else if (MODE == 2) {  // FFMA alone
#pragma unroll
                for (int i = 0; i < 8; ++i) x[i] = ffma(x[i], one, x[(i + 3) & 7]);
            } else if (MODE == 3) {  // IMAD + HFMA2 1:1  (> 2.0 => different sub-pipes)
#pragma unroll
                for (int i = 0; i < 8; i += 2) { x[i] = imad(x[i], b, x[(i + 3) & 7]); x[i + 1] = hfma_relu(x[i + 1], one, x[(i + 4) & 7]); }
            } else if (MODE == 4) {  // FFMA + HFMA2 1:1
#pragma unroll
                for (int i = 0; i < 8; i += 2) { x[i] = ffma(x[i], one, x[(i + 3) & 7]); x[i + 1] = hfma_relu(x[i + 1], one, x[(i + 4) & 7]); }
            } else if (MODE == 5) {  // FFMA + IMAD 1:1
#pragma unroll
                for (int i = 0; i < 8; i += 2) { x[i] = ffma(x[i], one, x[(i + 3) & 7]); x[i + 1] = imad(x[i + 1], b, x[(i + 4) & 7]); }
            } else if (MODE == 6) {  // HADD2 + HFMA2 1:1
#pragma unroll
                for (int i = 0; i < 8; i += 2) { x[i] = hadd2u(x[i], x[(i + 3) & 7]); x[i + 1] = hfma_relu(x[i + 1], one, x[(i + 4) & 7]); }
            } else if (MODE == 7) {  // IMAD + HFMA2 + VIMNMX(2 regs) 1:1:1 (9 instr)
#pragma unroll
                for (int i = 0; i < 9; i += 3) { x[i] = imad(x[i], b, x[(i + 4) % 9]); x[i + 1] = hfma_relu(x[i + 1], one, negulp); x[i + 2] = __vmaxs2(x[i + 2], x[(i + 5) % 9]); }
            } else if (MODE == 8) {  // IMAD + HFMA2 + VIMNMX3(3 regs) 1:1:1 (9 instr)
#pragma unroll
                for (int i = 0; i < 9; i += 3) { x[i] = imad(x[i], b, x[(i + 4) % 9]); x[i + 1] = hfma_relu(x[i + 1], one, negulp); x[i + 2] = __vimax3_s16x2(x[i + 2], x[(i + 5) % 9], x[(i + 7) % 9]); }
            } else if (MODE == 9) {  // cell with IMAD add: IMAD + VIMNMX3 + HFMA2.RELU, 4 chains (12 instr = 4 cells)
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    unsigned *y = x + 3 * c;
                    unsigned t = imad(b, a0, y[0]); unsigned m = __vimax3_s16x2(t, y[1], y[2]); y[2] = hfma_relu(m, one, negulp); y[0] = y[1]; y[1] = y[2];
                }
            } else if (MODE == 10) {  // LOP3 + VIMNMX 1:1 (both alu?)
#pragma unroll
                for (int i = 0; i < 8; i += 2) { x[i] = lop3(x[i], b, x[(i + 3) & 7]); x[i + 1] = __vmaxs2(x[i + 1], x[(i + 4) & 7]); }
            } else if (MODE == 11) {  // FADD + VIMNMX 1:1
#pragma unroll
                for (int i = 0; i < 8; i += 2) { x[i] = fadd(x[i], x[(i + 3) & 7]); x[i + 1] = __vmaxs2(x[i + 1], x[(i + 4) & 7]); }
            } else if (MODE == 12) {  // FADD alone
#pragma unroll
                for (int i = 0; i < 8; ++i) x[i] = fadd(x[i], x[(i + 3) & 7]);
            } else if (MODE == 13) {  // FADD + HFMA2 + VIMNMX 1:1:1
#pragma unroll
                for (int i = 0; i < 9; i += 3) { x[i] = fadd(x[i], x[(i + 4) % 9]); x[i + 1] = hfma_relu(x[i + 1], one, negulp); x[i + 2] = __vmaxs2(x[i + 2], x[(i + 5) % 9]); }
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
        run<0>("IMAD", 8, w, out, sms, ghz);
        run<1>("HFMA2.RELU", 8, w, out, sms, ghz);
        run<2>("FFMA", 8, w, out, sms, ghz);
        run<12>("FADD", 8, w, out, sms, ghz);
        run<3>("IMAD + HFMA2 1:1", 8, w, out, sms, ghz);
        run<4>("FFMA + HFMA2 1:1", 8, w, out, sms, ghz);
        run<5>("FFMA + IMAD 1:1", 8, w, out, sms, ghz);
        run<6>("HADD2 + HFMA2 1:1", 8, w, out, sms, ghz);
        run<10>("LOP3 + VIMNMX 1:1", 8, w, out, sms, ghz);
        run<11>("FADD + VIMNMX 1:1", 8, w, out, sms, ghz);
        run<7>("IMAD + HFMA2.RELU + VIMNMX(2r) 1:1:1", 9, w, out, sms, ghz);
        run<13>("FADD + HFMA2.RELU + VIMNMX(2r) 1:1:1", 9, w, out, sms, ghz);
        run<8>("IMAD + HFMA2.RELU + VIMNMX3(3r) 1:1:1", 9, w, out, sms, ghz);
        run<9>("cell: IMAD + VIMNMX3(3r) + HFMA2.RELU, 4 chains [12 instr = 4 cells]", 12, w, out, sms, ghz);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
