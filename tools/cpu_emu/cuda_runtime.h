// TEST INFRASTRUCTURE ONLY: a stand-in for <cuda_runtime.h> that lets g++ compile this repo's simpler CUDA kernels as
// plain C++ and run them on host threads (tools/cpu_emu/README.md).  One OS thread per CUDA thread of a block, blocks one
// after the other; __syncthreads / __syncwarp are C++20 barriers that exited threads drop out of.  It exists so that kernel
// logic and indexing can be checked (under AddressSanitizer too) where no GPU is available; it says nothing about
// performance, launch limits or device-only failure modes.
#pragma once
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <thread>
#include <vector>

#define __host__
#define __device__
#define __global__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))

struct uint3 { unsigned x, y, z; };
struct dim3 { unsigned x = 1, y = 1, z = 1; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(8) uint2 { unsigned x, y; };
struct alignas(16) uint4 { unsigned x, y, z, w; };
inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }

typedef int cudaError_t;
typedef void *cudaStream_t;
constexpr cudaError_t cudaSuccess = 0;
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount };
inline const char *cudaGetErrorString(cudaError_t) { return "emulated"; }
inline cudaError_t cudaMalloc(void **p, size_t n) { *p = std::malloc(n ? n : 1); return *p ? 0 : 2; }
inline cudaError_t cudaFree(void *p) { std::free(p); return 0; }
inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { std::memcpy(d, s, n); return 0; }
inline cudaError_t cudaSetDevice(int) { return 0; }
inline cudaError_t cudaGetLastError() { return 0; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
inline cudaError_t cudaDeviceSynchronize() { return 0; }
inline cudaError_t cudaDeviceGetAttribute(int *v, cudaDeviceAttr, int) { *v = 2; return 0; }  // a 2-SM "device": small passes
// what the C-ABI layer (api.cu) uses on top: everything synchronous, one "device"
typedef void *cudaEvent_t;
typedef void *cudaMemPool_t;
enum cudaMemPoolAttr { cudaMemPoolAttrReleaseThreshold };
constexpr unsigned cudaStreamNonBlocking = 1, cudaHostAllocDefault = 0;
inline cudaError_t cudaGetDeviceCount(int *n) { *n = 1; return 0; }
inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { std::memcpy(d, s, n); return 0; }
inline cudaError_t cudaMemset(void *d, int v, size_t n) { std::memset(d, v, n); return 0; }
inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t = nullptr) { std::memset(d, v, n); return 0; }
inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned) { *s = nullptr; return 0; }
inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
inline cudaError_t cudaHostAlloc(void **p, size_t n, unsigned) { *p = std::malloc(n ? n : 1); return *p ? 0 : 2; }
inline cudaError_t cudaFreeHost(void *p) { std::free(p); return 0; }
inline cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = nullptr; return 0; }
inline cudaError_t cudaEventDestroy(cudaEvent_t) { return 0; }
inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = nullptr) { return 0; }
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return 0; }
inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t, cudaEvent_t) { *ms = 0.0f; return 0; }
inline cudaError_t cudaDeviceGetDefaultMemPool(cudaMemPool_t *p, int) { *p = nullptr; return 0; }
inline cudaError_t cudaMemPoolSetAttribute(cudaMemPool_t, cudaMemPoolAttr, void *) { return 0; }
template <class T> inline cudaError_t cudaMalloc(T **p, size_t n) { return cudaMalloc(reinterpret_cast<void **>(p), n); }
template <class T> inline cudaError_t cudaHostAlloc(T **p, size_t n, unsigned f) { return cudaHostAlloc(reinterpret_cast<void **>(p), n, f); }

namespace emu {
struct Block {
    int live = 0;
    std::unique_ptr<std::barrier<>> all;
    std::vector<std::unique_ptr<std::barrier<>>> warp;
    std::vector<unsigned long long> slots;  // 32 exchange slots per warp (shuffles, votes)
    std::vector<unsigned char> dyn;         // dynamic shared memory of the block
};
inline thread_local uint3 t_threadIdx, t_blockIdx;
inline thread_local dim3 t_blockDim, t_gridDim;
inline thread_local Block *t_block = nullptr;

inline void *dyn_smem();
template <class F> void launch(dim3 grid, dim3 block, size_t smem, F &&body)
{
    const unsigned nthreads = block.x * block.y * block.z;
    for (unsigned bz = 0; bz < grid.z; ++bz)
        for (unsigned by = 0; by < grid.y; ++by)
            for (unsigned bx = 0; bx < grid.x; ++bx) {
                Block b;
                b.slots.assign(((nthreads + 31) / 32) * 32, 0ull);
                b.dyn.assign(smem + 16, 0);
                b.all = std::make_unique<std::barrier<>>((std::ptrdiff_t)nthreads);
                for (unsigned w = 0; w < (nthreads + 31) / 32; ++w)
                    b.warp.push_back(std::make_unique<std::barrier<>>((std::ptrdiff_t)std::min(32u, nthreads - 32 * w)));
                std::vector<std::thread> ts;
                ts.reserve(nthreads);
                for (unsigned t = 0; t < nthreads; ++t)
                    ts.emplace_back([&, t] {
                        t_threadIdx = {t % block.x, (t / block.x) % block.y, t / (block.x * block.y)};
                        t_blockIdx = {bx, by, bz};
                        t_blockDim = block; t_gridDim = grid; t_block = &b;
                        body();
                        b.slots[t] = 0;                    // ... nor in votes
                        b.all->arrive_and_drop();          // a thread that returned no longer takes part in barriers
                        b.warp[t / 32]->arrive_and_drop();
                    });
                for (auto &th : ts) th.join();
            }
}
inline void *dyn_smem() { return t_block->dyn.data(); }
inline unsigned flat_tid() { return t_threadIdx.x + t_threadIdx.y * t_blockDim.x + t_threadIdx.z * t_blockDim.x * t_blockDim.y; }
inline void warp_sync() { t_block->warp[flat_tid() / 32]->arrive_and_wait(); }
// every lane publishes a 64-bit value, then reads the lane it wants; lanes that already returned keep their last value
inline unsigned long long exchange(unsigned long long mine, int src_lane)
{
    const unsigned t = flat_tid(), base = (t / 32) * 32;
    t_block->slots[t] = mine;
    warp_sync();
    const unsigned long long r = (src_lane >= 0 && src_lane < 32 && base + (unsigned)src_lane < t_block->slots.size()) ? t_block->slots[base + src_lane] : mine;
    warp_sync();
    return r;
}
template <class T> inline T shfl(T v, int src_lane)
{
    static_assert(sizeof(T) <= 8, "shuffle of at most 64 bits");
    unsigned long long bits = 0;
    std::memcpy(&bits, &v, sizeof(T));
    bits = exchange(bits, src_lane);
    T r;
    std::memcpy(&r, &bits, sizeof(T));
    return r;
}
}  // namespace emu

#define threadIdx (emu::t_threadIdx)
#define blockIdx (emu::t_blockIdx)
#define blockDim (emu::t_blockDim)
#define gridDim (emu::t_gridDim)

inline void __syncthreads() { emu::t_block->all->arrive_and_wait(); }
inline void __syncwarp(unsigned = 0xffffffffu)
{
    const unsigned t = threadIdx.x + threadIdx.y * blockDim.x + threadIdx.z * blockDim.x * blockDim.y;
    emu::t_block->warp[t / 32]->arrive_and_wait();
}
// round-to-nearest single operations that must not be contracted: volatile keeps g++ from fusing them
inline float __fadd_rn(float a, float b) { volatile float r = a + b; return r; }
inline float __fsub_rn(float a, float b) { volatile float r = a - b; return r; }
inline float __fmul_rn(float a, float b) { volatile float r = a * b; return r; }
inline float __fsqrt_rn(float a) { return std::sqrt(a); }
inline float __fmaf_rn(float a, float b, float c) { return std::fmaf(a, b, c); }

// ---- warp collectives (every live lane of the warp must call them, as with a full mask on the device) ----
template <class T> inline T __shfl_sync(unsigned, T v, int src, int = 32) { return emu::shfl(v, src & 31); }
template <class T> inline T __shfl_xor_sync(unsigned, T v, int m, int = 32) { return emu::shfl(v, (int)((emu::flat_tid() & 31) ^ (unsigned)m)); }
template <class T> inline T __shfl_up_sync(unsigned, T v, unsigned d, int = 32)
{
    const int lane = (int)(emu::flat_tid() & 31);
    return emu::shfl(v, lane - (int)d >= 0 ? lane - (int)d : lane);
}
template <class T> inline T __shfl_down_sync(unsigned, T v, unsigned d, int = 32)
{
    const int lane = (int)(emu::flat_tid() & 31);
    return emu::shfl(v, lane + (int)d < 32 ? lane + (int)d : lane);
}
inline unsigned __ballot_sync(unsigned, int pred)
{
    const unsigned t = emu::flat_tid(), base = (t / 32) * 32;
    emu::t_block->slots[t] = pred ? 1ull : 0ull;
    emu::warp_sync();
    unsigned r = 0;
    for (unsigned l = 0; l < 32 && base + l < emu::t_block->slots.size(); ++l) r |= (unsigned)(emu::t_block->slots[base + l] & 1ull) << l;
    emu::warp_sync();
    emu::t_block->slots[t] = 0;  // a lane that returns later must not keep voting
    return r;
}
inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0; }
inline unsigned __match_any_sync(unsigned, int v)
{
    const unsigned t = emu::flat_tid(), base = (t / 32) * 32;
    emu::t_block->slots[t] = (unsigned long long)(unsigned)v | (1ull << 40);  // bit 40: this lane takes part
    emu::warp_sync();
    unsigned r = 0;
    for (unsigned l = 0; l < 32 && base + l < emu::t_block->slots.size(); ++l)
        if (emu::t_block->slots[base + l] == ((unsigned long long)(unsigned)v | (1ull << 40))) r |= 1u << l;
    emu::warp_sync();
    emu::t_block->slots[t] = 0;
    return r;
}
// warp reductions (REDUX on the device): every live lane contributes; lanes that already returned contribute their last value
template <class F> inline unsigned emu_reduce(unsigned v, F f)
{
    const unsigned t = emu::flat_tid(), base = (t / 32) * 32;
    emu::t_block->slots[t] = v;
    emu::warp_sync();
    unsigned r = v;
    for (unsigned l = 0; l < 32 && base + l < emu::t_block->slots.size(); ++l) r = f(r, (unsigned)emu::t_block->slots[base + l]);
    emu::warp_sync();
    emu::t_block->slots[t] = 0;
    return r;
}
inline unsigned __reduce_min_sync(unsigned, unsigned v) { return emu_reduce(v, [](unsigned a, unsigned b) { return a < b ? a : b; }); }
inline unsigned __reduce_max_sync(unsigned, unsigned v) { return emu_reduce(v, [](unsigned a, unsigned b) { return a > b ? a : b; }); }
inline unsigned __reduce_add_sync(unsigned, unsigned v)
{
    const unsigned t = emu::flat_tid(), base = (t / 32) * 32;
    emu::t_block->slots[t] = v;
    emu::warp_sync();
    unsigned r = 0;
    for (unsigned l = 0; l < 32 && base + l < emu::t_block->slots.size(); ++l) r += (unsigned)emu::t_block->slots[base + l];
    emu::warp_sync();
    emu::t_block->slots[t] = 0;
    return r;
}
inline int __all_sync(unsigned m, int pred) { return __ballot_sync(m, !pred) == 0; }
// ---- bit helpers ----
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }
inline int __ffs(int v) { return __builtin_ffs(v); }
inline int __clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
inline unsigned __float_as_uint(float f) { unsigned u; std::memcpy(&u, &f, 4); return u; }
inline float __uint_as_float(unsigned u) { float f; std::memcpy(&f, &u, 4); return f; }
inline int __float_as_int(float f) { int u; std::memcpy(&u, &f, 4); return u; }
inline float __int_as_float(int u) { float f; std::memcpy(&f, &u, 4); return f; }
template <class T> inline T __ldg(const T *p) { return *p; }
// ---- atomics (real ones: the emulated threads are OS threads) ----
template <class T> inline T atomicAdd(T *p, T v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
template <class T> inline T atomicOr(T *p, T v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
template <class T> inline T atomicExch(T *p, T v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
template <class T> inline T atomicCAS(T *p, T cmp, T v) { __atomic_compare_exchange_n(p, &cmp, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST); return cmp; }
template <class T> inline T atomicMin(T *p, T v)
{
    T old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    while (v < old && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
    return old;
}
template <class T> inline T atomicMax(T *p, T v)
{
    T old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    while (v > old && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
    return old;
}
inline void __threadfence() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
inline void __threadfence_system() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
// ---- launch attributes: accepted and ignored ----
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize };
template <class K> inline cudaError_t cudaFuncSetAttribute(K, cudaFuncAttribute, int) { return 0; }
template <class K> inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int *n, K, int, size_t) { *n = 1; return 0; }
using std::max;
using std::min;
