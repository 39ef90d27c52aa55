// Host bookkeeping of the L2 rerank stage: which candidates every query sees, exactly as the reference lays them out.
//   post_process_l2_dynamic  src/utils/post_processor.cpp:552-748   (find_sequences dynamic :74-201)
//   post_process_l2_static   src/utils/post_processor.cpp:1023-1162 (find_sequences static  :204-336)
// The reference expands the sparse ids of ALL queries into one flat list (ids whose position falls outside the bound are
// skipped, expansions at the start are clipped), de-duplicates the windows to embed, and cuts the flat list into
// per-query slices by a FIXED count per id - (2*stride - 1) in the dynamic variant, `stride` (1 on a dense index) in the
// static one.  That arithmetic is kept as it is; a slice that would run past the flat list (undefined behaviour in the
// reference) is refused.  CPU restatement: oracle/l2_oracle.py.
#include "common.h"

#include <algorithm>
#include <cstdlib>
#include <vector>

namespace drm {

namespace {

// the sparse branch of find_sequences: unique ids in order of first appearance + the flat mapping into them.
// The reference keeps position -> row in a std::unordered_map; the rows are numbered by first appearance, so any exact
// position index gives the same result.  An expansion touches 2 * stride - 1 CONSECUTIVE positions, so a direct table indexed by
// position (4 bytes per position below `bound`: 37 MB for E. coli, used up to 2^29 positions) walks memory sequentially where a
// hash scatters; above that an open-addressing table (linear probing, at most half full).  This loop is what is left of the
// stage's host time once the windows embed on the tensor cores.
int expand_flat(const uint64_t *ids, int64_t n_ids, uint64_t stride, uint64_t bound, L2Plan &plan)
{
    const uint64_t worst = (uint64_t)n_ids * (2 * stride - 1);  // every expanded position distinct
    if (!plan.direct_ids) plan.map.reserve((size_t)worst);
    if (plan.direct_ids) {  // the caller has every window's embedding: the flat list carries the window ids, nothing to index
        // where each id's expansion starts in the flat list (skipped ids add nothing, clipped ones less), then all ids in parallel
        std::vector<uint64_t> off((size_t)n_ids + 1);
        off[0] = 0;
        for (int64_t j = 0; j < n_ids; ++j) {
            const uint64_t p = ids[j] * stride;
            uint64_t cnt = 0;
            if (p < bound) cnt = std::min(p + stride, bound) - (p >= stride - 1 ? p - stride + 1 : 0);
            off[(size_t)j + 1] = off[(size_t)j] + cnt;
        }
        plan.map.resize((size_t)off[(size_t)n_ids]);
        uint32_t *out = plan.map.data();
#pragma omp parallel for schedule(static) if (n_ids > 65536)
        for (int64_t j = 0; j < n_ids; ++j) {
            const uint64_t p = ids[j] * stride;
            if (p >= bound) continue;
            const uint64_t start = p >= stride - 1 ? p - stride + 1 : 0;
            const uint64_t end = std::min(p + stride, bound);
            uint32_t *o = out + off[(size_t)j];
            for (uint64_t pos = start; pos < end; ++pos) *o++ = (uint32_t)pos;
        }
        return DRM_OK;
    }
    const char *tb = getenv("DRM_L2_PLAN_TABLE");  // "hash" forces the hashed table (A/B, tests)
    const bool direct = bound <= (1ull << 29) && !(tb && tb[0] == 'h');
    std::vector<uint32_t> row_of;          // direct: row + 1 of a position, 0 = not seen
    std::vector<uint64_t> keys;            // hashed: ~0 = empty (a position is below `bound` <= 2^64 - 1)
    std::vector<uint32_t> rows;
    uint64_t mask = 0;
    if (direct) row_of.assign((size_t)bound, 0u);
    else {
        const uint64_t most = std::min(worst, bound);  // distinct positions: all below `bound`
        uint64_t cap = 1024;
        while (cap < 2 * most && cap < (1ull << 33)) cap <<= 1;
        keys.assign((size_t)cap, ~0ull);
        rows.resize((size_t)cap);
        mask = cap - 1;
    }
    for (int64_t j = 0; j < n_ids; ++j) {
        const uint64_t p = ids[j] * stride;  // size_t arithmetic: wraps like the reference
        if (p >= bound) continue;
        const uint64_t start = p >= stride - 1 ? p - stride + 1 : 0;
        const uint64_t end = std::min(p + stride, bound);
        for (uint64_t pos = start; pos < end; ++pos) {
            uint32_t row;
            if (direct) {
                uint32_t &slot = row_of[(size_t)pos];
                if (slot == 0) {
                    if (plan.unique.size() >= 0xfffffffeull) return fail(DRM_EUNSUPPORTED, "more than 2^32 - 2 unique candidate windows");
                    plan.unique.push_back(pos);
                    slot = (uint32_t)plan.unique.size();
                }
                row = slot - 1;
            } else {
                uint64_t slot = ((pos * 0x9e3779b97f4a7c15ull) >> 20) & mask;
                while (keys[slot] != pos && keys[slot] != ~0ull) slot = (slot + 1) & mask;
                if (keys[slot] == ~0ull) {
                    if (plan.unique.size() >= 0xffffffffull) return fail(DRM_EUNSUPPORTED, "more than 2^32 - 1 unique candidate windows");
                    row = (uint32_t)plan.unique.size();
                    keys[slot] = pos;
                    rows[slot] = row;
                    plan.unique.push_back(pos);
                } else row = rows[slot];
            }
            plan.map.push_back(row);
        }
    }
    return DRM_OK;
}

}  // namespace

int l2_plan(int mode, const int64_t *neighbors, int64_t n, int64_t k_in, int64_t stride, int64_t k, int64_t k_clusters,
            uint64_t bound, L2Plan &plan, bool direct_ids)
{
    plan = L2Plan();
    plan.n = n;
    plan.direct_ids = direct_ids && bound < 0xffffffffull && !(mode != 1 && stride == 1);  // (the static 1-to-1 lookup has no index to skip)
    if (n <= 0 || k_in <= 0 || stride < 1 || k < 0 || k_clusters < 0) return fail(DRM_EINVAL, "bad arguments to the L2 rerank");
    int64_t per_slice;  // flat candidates the reference attributes to one query
    std::vector<uint64_t> ids;
    if (mode == 1) {
        if (k > k_clusters * 2 * stride && stride > 1)
            return fail(DRM_EKTOOLARGE, "Final k too large. Ensure k < k_clusters * 2 * stride to have enough candidates.");
        const int64_t per_q = std::min(stride == 1 ? k : k_clusters, k_in);
        plan.k_out = k;
        if (stride == 1) {  // "stride == 1 (dense), skipping reranker" (:638-664)
            plan.passthrough = true;
            plan.take = std::min(k, per_q);
            return DRM_OK;
        }
        ids.reserve((size_t)(n * per_q));
        for (int64_t q = 0; q < n; ++q)
            for (int64_t j = 0; j < per_q; ++j) ids.push_back((uint64_t)neighbors[q * k_in + j]);  // long -> size_t
        int rc = expand_flat(ids.data(), (int64_t)ids.size(), (uint64_t)stride, bound, plan);
        if (rc != DRM_OK) return rc;
        per_slice = per_q * (2 * stride - 1);
    } else {
        plan.k_out = k_clusters;  // the static variant hands k_clusters to batch_reranker (:1128-1135)
        ids.reserve((size_t)(n * k_in));
        for (int64_t j = 0; j < n * k_in; ++j) ids.push_back((uint64_t)neighbors[j]);
        if (stride == 1) {
            // 1-to-1 lookup that drops ids >= bound but records the ORIGINAL position as the mapping (:226-233)
            std::vector<uint32_t> keep;
            for (int64_t j = 0; j < (int64_t)ids.size(); ++j)
                if (ids[(size_t)j] < bound) { plan.unique.push_back(ids[(size_t)j]); keep.push_back((uint32_t)j); }
            for (uint32_t m : keep)
                if (m >= plan.unique.size()) return fail(DRM_EINVAL, "Invalid mapping index in expansion");
            plan.map = keep;
        } else {
            int rc = expand_flat(ids.data(), (int64_t)ids.size(), (uint64_t)stride, bound, plan);
            if (rc != DRM_OK) return rc;
        }
        per_slice = k_in * (stride > 1 ? stride : 1);
    }
    plan.starts.resize((size_t)n + 1);
    for (int64_t q = 0; q <= n; ++q) plan.starts[(size_t)q] = q * per_slice;
    if ((uint64_t)plan.starts[(size_t)n] > plan.map.size())
        return fail(DRM_EINVAL, "candidate slices run past the expanded list (an id was skipped or clipped: undefined in the reference)");
    return DRM_OK;
}

}  // namespace drm
