// Micro-benchmark: clocks per tcgen05.mma (cta_group::1, kind::f16, M = 128, K = 16) as a function of N and of the shared-memory
// operand layout - the K-major NO-SWIZZLE ("interleave") layout the embedder uses against K-major SWIZZLE_128B.  The operands
// hold zeros (only the issue / execution rate matters); one CTA per SM, one thread issues a chain of accumulating products
// and waits for the commit.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_umma tools/ubench_umma.cu
#include <cuda_runtime.h>

#include <cstdio>

__device__ __forceinline__ unsigned long long desc(unsigned saddr, unsigned lbo, unsigned sbo, unsigned layout)
{
    return (unsigned long long)((saddr >> 4) & 0x3fffu) | ((unsigned long long)((lbo >> 4) & 0x3fffu) << 16) |
           ((unsigned long long)((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46) | ((unsigned long long)layout << 61);
}

template <int N, int SWZ>
__global__ void __launch_bounds__(128, 1) umma_rate(int reps, int ksteps, unsigned long long *out)
{
    extern __shared__ __align__(1024) unsigned char sm[];
    __shared__ unsigned long long bar;
    __shared__ unsigned tslot;
    const int tid = threadIdx.x;
    for (int i = tid; i < 96 * 1024 / 16; i += 128) reinterpret_cast<uint4 *>(sm)[i] = make_uint4(0, 0, 0, 0);
    const unsigned bar_s = (unsigned)__cvta_generic_to_shared(&bar);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(&tslot)), "r"(256) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = tslot;
    const unsigned a_s = (unsigned)__cvta_generic_to_shared(sm), b_s = a_s + 32 * 1024;
    const unsigned idesc = (1u << 4) | ((unsigned)(N >> 3) << 17) | ((128u >> 4) << 24);
    long long t0 = 0, t1 = 0;
    if (tid == 0) {
        unsigned phase = 0;
        for (int r = -1; r < reps; ++r) {  // r = -1: warm-up
            if (r == 0) t0 = clock64();
            for (int ks = 0; ks < ksteps; ++ks) {
                unsigned long long ad, bd;
                if (SWZ) {  // rows of 128 bytes (K = 64 halves), 8-row groups 1024 bytes apart; a K step advances the start by 32 bytes
                    ad = desc(a_s + (ks & 3) * 32 + ((ks >> 2) & 1) * 16384, 16, 1024, 2);
                    bd = desc(b_s + (ks & 3) * 32 + ((ks >> 2) & 1) * (N * 128), 16, 1024, 2);
                } else {    // [k-group][row][16 B]: LBO = rows * 16, SBO = 128
                    ad = desc(a_s + (ks & 7) * 2 * 2048, 2048, 128, 0);   // operands re-used every 8 steps: 32 KB of A, <= 64 KB of B
                    bd = desc(b_s + (ks & 7) * 2 * (N * 16), N * 16, 128, 0);
                }
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(ad),
                             "l"(bd), "r"(idesc), "r"(ks ? 1u : 0u)
                             : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_s) : "memory");
            unsigned done = 0;
            do {
                asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}" : "=r"(done) : "r"(bar_s), "r"(phase) : "memory");
            } while (!done);
            phase ^= 1u;
        }
        t1 = clock64();
        if (blockIdx.x == 0) out[0] = (unsigned long long)(t1 - t0);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256) : "memory");
}

template <int N, int SWZ> void run(int ksteps, unsigned long long *d)
{
    const int reps = 200;
    cudaFuncSetAttribute(umma_rate<N, SWZ>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    umma_rate<N, SWZ><<<148, 128, 96 * 1024>>>(reps, ksteps, d);
    unsigned long long h = 0;
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    const double per = (double)h / reps / ksteps;
    printf("N = %3d  %-12s  %2d products per commit: %7.1f clocks per product  (%5.1f %% of 4096 multiply-adds per clock: ideal %5.1f)  %s\n", N, SWZ ? "swizzle 128B" : "no swizzle",
           ksteps, per, 100.0 * (128.0 * N * 16 / 4096.0) / per, 128.0 * N * 16 / 4096.0, e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main()
{
    unsigned long long *d;
    cudaMalloc(&d, 8);
    for (int ks : {8, 16, 32, 64}) {
        run<64, 0>(ks, d); run<128, 0>(ks, d); run<192, 0>(ks, d); run<256, 0>(ks, d);
        run<64, 1>(ks, d); run<128, 1>(ks, d); run<192, 1>(ks, d); run<256, 1>(ks, d);
    }
    return 0;
}
