// L2 rerank on the device: batch_reranker (src/utils/reranker.cpp:98-196) over candidate embeddings produced by the
// embedder (embed_kernel.cu).  One warp per query: the lanes compute calc_l2_dist (src/utils/metrics.cpp:47-61) for the
// query's slice of the flat candidate list, then lane 0 runs libstdc++'s partial_sort (heap select + sort_heap,
// bits/stl_algo.h / stl_heap.h) on the slice - sequential, because the order in which equal distances come out
// (identical windows at different positions of a repetitive reference) is decided by that algorithm's data movement.
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
__host__ __device__ inline void l2_push_heap(long long *first, long long hole, long long top, long long value, const float *dist)
{
    long long parent = (hole - 1) / 2;
    while (hole > top && dist[first[parent]] < dist[value]) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}

__host__ __device__ inline void l2_adjust_heap(long long *first, long long hole, long long len, long long value, const float *dist)
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

__host__ __device__ inline void l2_partial_sort(const float *dist, long long n, long long k, long long *idx)
{
    for (long long i = 0; i < n; ++i) idx[i] = i;  // std::iota
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
                const long long value = idx[i];
                idx[i] = idx[0];
                l2_adjust_heap(idx, 0, k, value, dist);
            }
        }
    }
    long long last = k;  // __sort_heap(first, middle)
    while (last > 1) {
        --last;
        const long long value = idx[last];
        idx[last] = idx[0];
        l2_adjust_heap(idx, 0, last, value, dist);
    }
}

__global__ void __launch_bounds__(L2_WARPS * 32)
l2_rerank_kernel(const float *__restrict__ emb, const uint32_t *__restrict__ map, const long long *__restrict__ starts,
                 const float *__restrict__ q, const long long n, const int k, float *dist_scr, long long *idx_scr,
                 float *__restrict__ out_dist, uint32_t *__restrict__ out_rows, long long *__restrict__ counts)
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
    const float *qe = q + (size_t)query * EMB_D;
    for (long long c = lane; c < nc; c += 32) {
        const float *e = emb + (size_t)map[a + c] * EMB_D;
        float sum = 0.0f;
#pragma unroll 8
        for (int i = 0; i < EMB_D; ++i) {
            const float diff = __fsub_rn(e[i], qe[i]);
            sum = __fadd_rn(sum, __fmul_rn(diff, diff));
        }
        dist_scr[a + c] = __fsqrt_rn(sum);
    }
    __syncwarp(FULL);
    if (lane != 0) return;
    const float *dist = dist_scr + a;
    long long *idx = idx_scr + a;
    l2_partial_sort(dist, nc, k, idx);
    for (int i = 0; i < k; ++i) {
        out_dist[(size_t)query * k + i] = dist[idx[i]];
        out_rows[(size_t)query * k + i] = map[a + idx[i]];
    }
    counts[query] = k;
}

}  // namespace

void l2_select_host(const float *dist, int64_t n, int64_t k, int64_t *idx)
{
    l2_partial_sort(dist, (long long)n, (long long)k, reinterpret_cast<long long *>(idx));
}

int launch_l2_rerank(const float *emb, const uint32_t *map, const int64_t *starts, const float *q, int64_t n, int k,
                     float *dist_scr, int64_t *idx_scr, float *out_dist, uint32_t *out_rows, int64_t *counts,
                     cudaStream_t stream)
{
    if (n <= 0) return DRM_OK;
    if (k < 1) return fail(DRM_EINVAL, "k must be >= 1");
    const unsigned blocks = (unsigned)((n + L2_WARPS - 1) / L2_WARPS);
    l2_rerank_kernel<<<blocks, L2_WARPS * 32, 0, stream>>>(emb, map, reinterpret_cast<const long long *>(starts), q, (long long)n, k, dist_scr,
                                                          reinterpret_cast<long long *>(idx_scr), out_dist, out_rows,
                                                          reinterpret_cast<long long *>(counts));
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

}  // namespace drm
