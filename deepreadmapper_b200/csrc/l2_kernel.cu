// L2 rerank on the device: batch_reranker (src/utils/reranker.cpp:98-196) over candidate embeddings produced by the
// embedder (embed_kernel.cu).  Two kernels:
//   l2_dist_kernel    one warp per query; the lanes compute calc_l2_dist (src/utils/metrics.cpp:47-61) for the query's slice
//                     of the flat candidate list (each lane streams candidate rows; 512 B per row);
//   l2_select_kernel  one THREAD per query runs libstdc++'s partial_sort (heap select + sort_heap, bits/stl_algo.h /
//                     stl_heap.h) on its slice.  The algorithm stays sequential per query because the order in which equal
//                     distances come out (identical windows at different positions of a repetitive reference) is decided by its
//                     data movement - but 32 queries now run per warp instead of one (round 1 ran it on lane 0 of a warp: 1/32 of
//                     the lanes), on 32-bit indices.
// CPU restatement: oracle/l2_oracle.c.
//
// Distance arithmetic: g++ compiles the reference's `sum += diff * diff` loop to vmulps + ordered vaddss - products are
// rounded before they are added, no FMA (oracle/l2_oracle.c header, tests/test_oracle_l2.py) - so the chain here is
// __fsub_rn / __fmul_rn / __fadd_rn in index order and __fsqrt_rn at the end: bit-equal distances.
#include "common.h"

namespace drm {

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int L2_WARPS = 4;
constexpr int EMB_D = 128;

// ---- std::partial_sort(idx, idx + k, idx + n, [&](a, b) { return dist[a] < dist[b]; }) ----
template <class IT>
__host__ __device__ inline void l2_push_heap(IT *first, long long hole, long long top, IT value, const float *dist)
{
    long long parent = (hole - 1) / 2;
    while (hole > top && dist[first[parent]] < dist[value]) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}

template <class IT>
__host__ __device__ inline void l2_adjust_heap(IT *first, long long hole, long long len, IT value, const float *dist)
{
    const long long top = hole;
    long long child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (dist[first[child]] < dist[first[child - 1]]) child--;
        first[hole] = first[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        first[hole] = first[child - 1];
        hole = child - 1;
    }
    l2_push_heap(first, hole, top, value, dist);
}

template <class IT>
__host__ __device__ inline void l2_partial_sort(const float *dist, long long n, long long k, IT *idx)
{
    for (long long i = 0; i < n; ++i) idx[i] = (IT)i;  // std::iota
    if (k > n) k = n;
    if (k >= 2) {  // __make_heap(first, middle)
        long long parent = (k - 2) / 2;
        for (;;) {
            l2_adjust_heap(idx, parent, k, idx[parent], dist);
            if (parent == 0) break;
            parent--;
        }
    }
    if (k >= 1) {
        for (long long i = k; i < n; ++i) {
            if (dist[idx[i]] < dist[idx[0]]) {  // __pop_heap(first, middle, i)
                const IT value = idx[i];
                idx[i] = idx[0];
                l2_adjust_heap(idx, 0, k, value, dist);
            }
        }
    }
    long long last = k;  // __sort_heap(first, middle)
    while (last > 1) {
        --last;
        const IT value = idx[last];
        idx[last] = idx[0];
        l2_adjust_heap(idx, 0, last, value, dist);
    }
}

__global__ void __launch_bounds__(L2_WARPS * 32)
l2_dist_kernel(const float *__restrict__ emb, const uint32_t *__restrict__ map, const long long *__restrict__ starts,
               const float *__restrict__ q, const long long n, const int k, float *__restrict__ dist_scr, long long *__restrict__ counts)
{
    const int lane = threadIdx.x & 31;
    const long long query = (long long)blockIdx.x * L2_WARPS + (threadIdx.x >> 5);
    if (query >= n) return;
    const long long a = starts[query], nc = starts[query + 1] - a;
    if (nc == 0) {
        if (lane == 0) counts[query] = 0;
        return;
    }
    if (nc < k) {  // "Not enough candidates (nc < k) for query q"
        if (lane == 0) counts[query] = -nc - 1;
        return;
    }
    if (lane == 0) counts[query] = k;
    const float4 *qe = reinterpret_cast<const float4 *>(q + (size_t)query * EMB_D);
    for (long long c = lane; c < nc; c += 32) {
        const float4 *e = reinterpret_cast<const float4 *>(emb + (size_t)map[a + c] * EMB_D);
        float sum = 0.0f;
#pragma unroll 4
        for (int i = 0; i < EMB_D / 4; ++i) {
            const float4 ev = __ldg(e + i), qv = __ldg(qe + i);
            float d0 = __fsub_rn(ev.x, qv.x); sum = __fadd_rn(sum, __fmul_rn(d0, d0));
            float d1 = __fsub_rn(ev.y, qv.y); sum = __fadd_rn(sum, __fmul_rn(d1, d1));
            float d2 = __fsub_rn(ev.z, qv.z); sum = __fadd_rn(sum, __fmul_rn(d2, d2));
            float d3 = __fsub_rn(ev.w, qv.w); sum = __fadd_rn(sum, __fmul_rn(d3, d3));
        }
        dist_scr[a + c] = __fsqrt_rn(sum);
    }
}

__global__ void __launch_bounds__(128)
l2_select_kernel(const uint32_t *__restrict__ map, const long long *__restrict__ starts, const long long n, const int k,
                 const float *__restrict__ dist_scr, int *__restrict__ idx_scr, float *__restrict__ out_dist,
                 uint32_t *__restrict__ out_rows, const long long *__restrict__ counts)
{
    const long long query = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (query >= n || counts[query] != k) return;
    const long long a = starts[query], nc = starts[query + 1] - a;
    const float *dist = dist_scr + a;
    int *idx = idx_scr + a;
    l2_partial_sort<int>(dist, nc, k, idx);
    for (int i = 0; i < k; ++i) {
        out_dist[(size_t)query * k + i] = dist[idx[i]];
        out_rows[(size_t)query * k + i] = map[a + idx[i]];
    }
}

}  // namespace

void l2_select_host(const float *dist, int64_t n, int64_t k, int64_t *idx)
{
    l2_partial_sort<long long>(dist, (long long)n, (long long)k, reinterpret_cast<long long *>(idx));
}

int launch_l2_rerank(const float *emb, const uint32_t *map, const int64_t *starts, const float *q, int64_t n, int k,
                     float *dist_scr, int64_t *idx_scr, float *out_dist, uint32_t *out_rows, int64_t *counts,
                     cudaStream_t stream)
{
    if (n <= 0) return DRM_OK;
    if (k < 1) return fail(DRM_EINVAL, "k must be >= 1");
    const unsigned blocks = (unsigned)((n + L2_WARPS - 1) / L2_WARPS);
    l2_dist_kernel<<<blocks, L2_WARPS * 32, 0, stream>>>(emb, map, reinterpret_cast<const long long *>(starts), q, (long long)n, k, dist_scr,
                                                        reinterpret_cast<long long *>(counts));
    DRM_CUDA_TRY(cudaGetLastError());
    l2_select_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(map, reinterpret_cast<const long long *>(starts), (long long)n, k, dist_scr,
                                                                     reinterpret_cast<int *>(idx_scr), out_dist, out_rows,
                                                                     reinterpret_cast<const long long *>(counts));
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

}  // namespace drm
