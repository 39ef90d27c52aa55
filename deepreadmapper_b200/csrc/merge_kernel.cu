// Per-read K-way merge for a position-sharded index (BASELINE.json config 5): after the all-gather of
// every shard's local top-k, one warp per read merges the n_shards ascending lists into the global
// top-k ordered by (distance, id).  Lane s walks shard s; each output is a warp min-reduction.
#include "common.h"

#include <cfloat>

namespace drm {

namespace {
typedef unsigned long long u64;
constexpr unsigned FULL = 0xffffffffu;

struct Key { unsigned d; u64 id; };
__device__ __forceinline__ bool key_less(unsigned d1, u64 i1, unsigned d2, u64 i2) { return d1 < d2 || (d1 == d2 && i1 < i2); }

__global__ void merge_topk_kernel(const float *__restrict__ D_all, const long long *__restrict__ I_all, int n_shards,
                                  long long n, int k, float *__restrict__ D_out, long long *__restrict__ I_out, long long shard_stride)
{
    const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= n) return;
    const float *Ds = D_all + ((size_t)lane * shard_stride + r) * k;
    const long long *Is = I_all + ((size_t)lane * shard_stride + r) * k;
    int head = 0;
    const bool owner = lane < n_shards;
    // unfilled slots (-1) compare as the largest id, so they sort after every real entry
    unsigned hd = 0xffffffffu;
    u64 hi = ~0ull;
    if (owner && k > 0) { hd = __float_as_uint(Ds[0]); hi = (u64)Is[0]; }
    for (int o = 0; o < k; ++o) {
        unsigned bd = hd; u64 bi = hi; int bl = lane;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            unsigned od = __shfl_xor_sync(FULL, bd, s);
            u64 oi = __shfl_xor_sync(FULL, bi, s);
            int ol = __shfl_xor_sync(FULL, bl, s);
            if (key_less(od, oi, bd, bi) || (od == bd && oi == bi && ol < bl)) { bd = od; bi = oi; bl = ol; }
        }
        if (lane == 0) {
            D_out[r * k + o] = bd == 0xffffffffu ? FLT_MAX : __uint_as_float(bd);
            I_out[r * k + o] = bd == 0xffffffffu ? -1 : (long long)bi;
        }
        if (lane == bl && owner) {
            ++head;
            if (head < k) { hd = __float_as_uint(Ds[head]); hi = (u64)Is[head]; }
            else { hd = 0xffffffffu; hi = ~0ull; }
        }
    }
}
}  // namespace

int launch_merge_topk(const float *D_all, const int64_t *I_all, int n_shards, int64_t n, int k, float *D_out,
                      int64_t *I_out, cudaStream_t stream, int64_t shard_stride)
{
    if (n <= 0 || k <= 0) return DRM_OK;
    if (n_shards < 1 || n_shards > 32) return fail(DRM_EINVAL, "merge supports 1..32 shards");
    const int warps_per_cta = 8;
    long long ctas = (n + warps_per_cta - 1) / warps_per_cta;
    merge_topk_kernel<<<(unsigned)ctas, warps_per_cta * 32, 0, stream>>>(D_all, reinterpret_cast<const long long *>(I_all), n_shards, n, k,
                                                                       D_out, reinterpret_cast<long long *>(I_out), shard_stride < 0 ? n : shard_stride);
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

}  // namespace drm
