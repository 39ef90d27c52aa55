/* TEST INFRASTRUCTURE ONLY - CPU restatement of the numeric core of the reference's L2 rerank (the
 * "live" rerank variants the embedder enables).  Never linked into the product.
 *
 *   orc_l2_dist            calc_l2_dist            src/utils/metrics.cpp:47-61
 *   orc_partial_sort_asc   std::partial_sort(idx, idx+k, idx+n, [&](a,b){ return d[a] < d[b]; }) as libstdc++'s
 *                          bits/stl_algo.h __heap_select + __sort_heap do it (src/utils/reranker.cpp:86-88,165-167)
 *   orc_batch_rerank_one   one query of batch_reranker  src/utils/reranker.cpp:121-183
 * The flattening / expansion / boundary logic around it lives in oracle/l2_oracle.py.
 *
 * calc_l2_dist is `sum += diff * diff` in a loop.  g++ -O3 (x86-64-v3 in oracle/_ref, -march=native in the reference's
 * own build) vectorises the subtraction and the multiplication (vsubps / vmulps) and keeps the additions as an ordered
 * scalar chain (vaddss), so the products are rounded before they are added: the UNFUSED chain (use_fma = 0) is what the
 * compiled reference computes, bit for bit (tests/test_oracle_l2.py); use_fma = 1 is kept to show the difference. */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

float orc_l2_dist(const float *a, const float *b, int64_t d, int use_fma)
{
    volatile float sum = 0.0f; /* volatile: keep the compiler from re-associating or vectorising the chain */
    for (int64_t i = 0; i < d; ++i) {
        const float diff = a[i] - b[i];
        if (use_fma) sum = fmaf(diff, diff, sum);
        else {
            volatile float sq = diff * diff;
            sum = sum + sq;
        }
    }
    return sqrtf(sum);
}

#define L2_COMP(a, b) (dist[(a)] < dist[(b)])

static void l2_push_heap(int64_t *first, int64_t hole, int64_t top, int64_t value, const float *dist)
{
    int64_t parent = (hole - 1) / 2;
    while (hole > top && L2_COMP(first[parent], value)) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}

static void l2_adjust_heap(int64_t *first, int64_t hole, int64_t len, int64_t value, const float *dist)
{
    const int64_t top = hole;
    int64_t child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (L2_COMP(first[child], first[child - 1])) child--;
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

/* idx[0..n) = iota, permuted in place; idx[0..k) is what the reference gathers from (ascending distance). */
void orc_partial_sort_asc(const float *dist, int64_t n, int64_t k, int64_t *idx)
{
    for (int64_t i = 0; i < n; ++i) idx[i] = i;
    if (k > n) k = n;
    if (k >= 2) {
        int64_t parent = (k - 2) / 2;
        for (;;) {
            l2_adjust_heap(idx, parent, k, idx[parent], dist);
            if (parent == 0) break;
            parent--;
        }
    }
    if (k >= 1) {
        for (int64_t i = k; i < n; ++i) {
            if (L2_COMP(idx[i], idx[0])) {
                const int64_t value = idx[i];
                idx[i] = idx[0];
                l2_adjust_heap(idx, 0, k, value, dist);
            }
        }
    }
    int64_t last = k;
    while (last > 1) {
        --last;
        const int64_t value = idx[last];
        idx[last] = idx[0];
        l2_adjust_heap(idx, 0, last, value, dist);
    }
}

/* One query of batch_reranker: candidates c = 0..n_cands-1 have embedding rows emb_ids[c] of `emb` (rows of d floats)
 * and dense ids dense_ids[c].  Returns k, 0 for an empty candidate list, or -1 when n_cands < k (the reference throws
 * "Not enough candidates").  Outputs k distances and k dense ids. */
int64_t orc_batch_rerank_one(const float *emb, const uint64_t *emb_ids, const uint64_t *dense_ids, int64_t n_cands,
                             const float *query, int64_t d, int64_t k, int use_fma, float *out_dist, uint64_t *out_ids)
{
    if (n_cands == 0) return 0;
    if (n_cands < k) return -1;
    float *dist = (float *)malloc(sizeof(float) * (size_t)n_cands);
    int64_t *idx = (int64_t *)malloc(sizeof(int64_t) * (size_t)n_cands);
    for (int64_t c = 0; c < n_cands; ++c) dist[c] = orc_l2_dist(emb + emb_ids[c] * (uint64_t)d, query, d, use_fma);
    orc_partial_sort_asc(dist, n_cands, k, idx);
    for (int64_t i = 0; i < k; ++i) { out_dist[i] = dist[idx[i]]; out_ids[i] = dense_ids[idx[i]]; }
    free(idx);
    free(dist);
    return k;
}
