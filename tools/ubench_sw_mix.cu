// Micro-benchmark #2: which pipes can carry the Smith-Waterman cell recurrence on sm_100a.
// Each mode is an unrolled loop of independent chains; the output is warp-instructions per clock per SM
// (4.0 = the issue limit of the 4 SMSPs) and, for the cell mixes, cell PAIRS (one s16x2 register) per clock per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_sw_mix tools/ubench_sw_mix.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>

#define ITER 2048
#define U8(OP) OP(0) OP(1) OP(2) OP(3) OP(4) OP(5) OP(6) OP(7)

template <int MODE> __global__ void __launch_bounds__(512) k(unsigned *out, unsigned a0, unsigned b0, unsigned one, const uint4 *gtab)
{
    __shared__ uint4 tab[25 * 10 * 8];  // [ab][4-column group][8 bank-group slots]
    for (int i = threadIdx.x; i < 25 * 10 * 8; i += blockDim.x) tab[i] = gtab[i];
    __syncthreads();
    unsigned x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = threadIdx.x * (2 * i + 1) + a0 + i;
    const unsigned m1 = 0xffffffffu, b = b0;
    const int slot = threadIdx.x & 7;
    unsigned ab = (threadIdx.x * 7u + a0) % 25u;
#pragma unroll 1
    for (int it = 0; it < ITER; ++it) {
        if (MODE == 0) {  // LOP3 (a & b) ^ c
#define OP(i) x[i] = (x[i] & x[(i + 1) & 7]) ^ b;
            U8(OP) U8(OP) U8(OP) U8(OP)
#undef OP
        } else if (MODE == 1) {  // IMAD with a runtime multiplier
#define OP(i) x[i] = x[i] * one + x[(i + 1) & 7];
            U8(OP) U8(OP) U8(OP) U8(OP)
#undef OP
        } else if (MODE == 2) {  // HMNMX2
#define OP(i) { __half2 r = __hmax2(*reinterpret_cast<__half2 *>(&x[i]), *reinterpret_cast<__half2 *>(&x[(i + 1) & 7])); x[i] = *reinterpret_cast<unsigned *>(&r); }
            U8(OP) U8(OP) U8(OP) U8(OP)
#undef OP
        } else if (MODE == 3) {  // HADD2
#define OP(i) { __half2 r = __hadd2(*reinterpret_cast<__half2 *>(&x[i]), *reinterpret_cast<__half2 *>(&x[(i + 1) & 7])); x[i] = *reinterpret_cast<unsigned *>(&r); }
            U8(OP) U8(OP) U8(OP) U8(OP)
#undef OP
        } else if (MODE == 4) {  // HSET2.EQ (-> 1.0 / 0.0 per half)
#define OP(i) { __half2 r = __heq2(*reinterpret_cast<__half2 *>(&x[i]), *reinterpret_cast<__half2 *>(&x[(i + 1) & 7])); x[i] ^= *reinterpret_cast<unsigned *>(&r); }
            U8(OP) U8(OP) U8(OP) U8(OP)
#undef OP
        } else if (MODE == 5) {  // cell mix A (shipping kernel): PRMT + IADD + VIMNMX3 + VIADDMNMX.RELU (+ half a VIMNMX3)
#define CELL(d, u, l, q, s) { unsigned s2 = __byte_perm(q, 0u, s); unsigned t = d + s2; unsigned m = __vimax3_s16x2(t, u, l); l = __viaddmax_s16x2_relu(m, m1, 0u); d = u; u = l; best = __vimax3_s16x2(best, m, t); }
            unsigned best = x[6];
#pragma unroll
            for (int r = 0; r < 8; ++r) { CELL(x[0], x[1], x[2], x[3], x[7]) CELL(x[4], x[5], x[2], x[3], x[7]) }
            x[6] = best;
#undef CELL
        } else if (MODE == 6) {  // cell mix B: IMAD(one-hot product) + LOP3 mask + IMAD add + VIMNMX3 + VIADDMNMX.RELU
#define CELL(d, u, l, q, s) { unsigned p = q * s; unsigned s2 = p & 0x00400040u; unsigned t = s2 * one + d; unsigned m = __vimax3_s16x2(t, u, l); l = __viaddmax_s16x2_relu(m, 0xffe0ffe0u, 0u); d = u; u = l; best = __vimax3_s16x2(best, m, t); }
            unsigned best = x[6];
#pragma unroll
            for (int r = 0; r < 8; ++r) { CELL(x[0], x[1], x[2], x[3], x[7]) CELL(x[4], x[5], x[2], x[3], x[7]) }
            x[6] = best;
#undef CELL
        } else if (MODE == 7 || MODE == 8) {  // cell mix C: LDS.128 profile (4 columns per load) + add + VIMNMX3 + VIADDMNMX.RELU
            unsigned best = x[6];
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const uint4 pr = tab[(ab * 10 + g) * 8 + slot];
                const unsigned s4[4] = {pr.x, pr.y, pr.z, pr.w};
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    unsigned t = MODE == 7 ? s4[c] * one + x[0] : s4[c] + x[0];
                    unsigned m = __vimax3_s16x2(t, x[1], x[2]);
                    x[2] = __viaddmax_s16x2_relu(m, m1, 0u);
                    x[0] = x[1]; x[1] = x[2];
                    if (c & 1) best = __vimax3_s16x2(best, m, x[3]);
                    x[3] = m;
                }
            }
            x[6] = best;
            ab = (ab + 7u + (x[2] & 1u)) % 25u;
        }
    }
    unsigned r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r ^ ab;
}

template <int MODE> void run(const char *name, double instr_per_iter, double pairs_per_iter, unsigned *out, const uint4 *tab, int sms, double ghz)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int ctas = sms * 4, threads = 512;
    k<MODE><<<ctas, threads>>>(out, 1, 3, 1, tab);
    cudaEventRecord(a);
    k<MODE><<<ctas, threads>>>(out, 1, 3, 1, tab);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    const double warps = (double)ctas * threads / 32;
    const double clk = ms * 1e-3 * ghz * 1e9;
    printf("%-66s %8.3f ms  %5.2f warp-instr/clk/SM", name, ms, warps * ITER * instr_per_iter / clk / sms);
    if (pairs_per_iter > 0) printf("  %6.2f cell-pairs/clk/SM = %6.2f TCUPS at %.3f GHz", warps * 32 * ITER * pairs_per_iter / clk / sms, warps * 32 * ITER * pairs_per_iter * 2 / (ms * 1e-3) / 1e12, ghz);
    printf("\n");
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ghz = khz / 1e6;
    printf("%s, %d SMs, max clock %.3f GHz (rates below assume the max clock)\n", p.name, p.multiProcessorCount, ghz);
    unsigned *out; uint4 *tab;
    cudaMalloc(&out, (size_t)p.multiProcessorCount * 4 * 512 * 4);
    cudaMalloc(&tab, 25 * 10 * 8 * sizeof(uint4));
    cudaMemset(tab, 0, 25 * 10 * 8 * sizeof(uint4));
    const int sms = p.multiProcessorCount;
    run<0>("LOP3", 32, 0, out, tab, sms, ghz);
    run<1>("IMAD (runtime multiplier)", 32, 0, out, tab, sms, ghz);
    run<2>("HMNMX2", 32, 0, out, tab, sms, ghz);
    run<3>("HADD2", 32, 0, out, tab, sms, ghz);
    run<4>("HSET2.EQ + LOP", 64, 0, out, tab, sms, ghz);
    run<5>("cell mix A: PRMT + IADD + VIMNMX3 + VIADDMNMX + VIMNMX3(best)", 16 * 5, 16, out, tab, sms, ghz);
    run<6>("cell mix B: IMAD + LOP3 + IMAD + VIMNMX3 + VIADDMNMX + VIMNMX3", 16 * 6, 16, out, tab, sms, ghz);
    run<7>("cell mix C: LDS.128/4 + IMAD + VIMNMX3 + VIADDMNMX + VIMNMX3/2", 16 * 3.75, 16, out, tab, sms, ghz);
    run<8>("cell mix C': LDS.128/4 + IADD + VIMNMX3 + VIADDMNMX + VIMNMX3/2", 16 * 3.75, 16, out, tab, sms, ghz);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
