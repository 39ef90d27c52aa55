// CPU index builder behind the hnswpq_index CLI: product-quantizer training + encoding and HNSW
// construction with ADC / symmetric PQ distances, written to the FAISS "IHNp" format.
//
// Mirrors what the reference asks FAISS to do in build_faiss_index (src/hnswpq/index.cpp:86-193):
//   IndexHNSWPQ(dim, M_pq, M_hnsw, nbits); hnsw.efConstruction = EFC; train(n_train, sample); add(...);
// i.e. FAISS ProductQuantizer::train (k-means per sub-space, at most 256 points per centroid),
// IndexPQ::add (nearest-centroid codes) and hnsw_add_vertices / HNSW::add_with_locks (level law
// 1/ln(M), insertion by descending level, efConstruction beam, shrink_neighbor_list heuristic on
// symmetric distances).  It is not on the timed search path; it exists because the path needs an index
// file and FAISS is not available.  Graph construction is multi-threaded and, like FAISS's, not
// run-to-run reproducible.
#include "common.h"

#include <omp.h>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <numeric>
#include <queue>
#include <random>

namespace drm {

namespace {

// ---------------------------------------------------------------- k-means for one sub-quantizer
void kmeans(const float *x, int64_t n, int dsub, int k, int iters, uint32_t seed, float *cent)
{
    std::vector<int64_t> start;
    builder_kmeans_init_rows(n, k, seed, start);
    for (int c = 0; c < k; ++c) memcpy(cent + (size_t)c * dsub, x + (size_t)start[(size_t)c] * dsub, sizeof(float) * dsub);
    std::vector<int> assign((size_t)n);
    std::vector<double> sum((size_t)k * dsub);
    std::vector<int64_t> cnt((size_t)k);
    for (int it = 0; it < iters; ++it) {
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            const float *xi = x + (size_t)i * dsub;
            float best = FLT_MAX; int bi = 0;
            for (int c = 0; c < k; ++c) {
                const float *cc = cent + (size_t)c * dsub;
                float s = 0;
                for (int t = 0; t < dsub; ++t) { float e = xi[t] - cc[t]; s += e * e; }
                if (s < best) { best = s; bi = c; }
            }
            assign[(size_t)i] = bi;
        }
        std::fill(sum.begin(), sum.end(), 0.0);
        std::fill(cnt.begin(), cnt.end(), 0);
        for (int64_t i = 0; i < n; ++i) {
            int c = assign[(size_t)i];
            cnt[(size_t)c]++;
            for (int t = 0; t < dsub; ++t) sum[(size_t)c * dsub + t] += x[(size_t)i * dsub + t];
        }
        builder_kmeans_update(sum.data(), cnt.data(), k, dsub, cent);
    }
}

struct Graph {
    HostIndex &ix;
    int M;
    std::vector<float> sdc;  // [M_pq][256][256]
    std::vector<omp_lock_t> locks;
    explicit Graph(HostIndex &i) : ix(i), M(i.M_hnsw()) {}
    int32_t *row(int64_t v, int level, int &width)
    {
        width = ix.cum[(size_t)level + 1] - ix.cum[(size_t)level];
        return ix.neighbors.data() + ix.offsets[(size_t)v] + ix.cum[(size_t)level];
    }
    float sym(int64_t a, int64_t b) const
    {
        const uint8_t *ca = &ix.codes[(size_t)a * ix.code_size], *cb = &ix.codes[(size_t)b * ix.code_size];
        float s = 0;
        for (uint64_t m = 0; m < ix.M; ++m) s += sdc[(m * 256 + ca[m]) * 256 + cb[m]];
        return s;
    }
};

struct Cand { float d; int32_t id; };
struct Farther { bool operator()(const Cand &a, const Cand &b) const { return a.d < b.d; } };   // max-heap on d
struct Closer { bool operator()(const Cand &a, const Cand &b) const { return a.d > b.d; } };    // min-heap on d

// HNSW::shrink_neighbor_list: input sorted ascending by distance to the query vertex
void shrink(const Graph &g, std::vector<Cand> &sorted, size_t max_size, std::vector<Cand> &out)
{
    out.clear();
    for (const Cand &v1 : sorted) {
        bool good = true;
        for (const Cand &v2 : out)
            if (g.sym(v1.id, v2.id) < v1.d) { good = false; break; }
        if (good) {
            out.push_back(v1);
            if (out.size() >= max_size) return;
        }
    }
}

struct Worker {
    Graph &g;
    std::vector<uint16_t> visited;
    uint16_t stamp = 0;
    std::vector<float> T;
    std::vector<Cand> tmp, sel;
    explicit Worker(Graph &gr) : g(gr), visited((size_t)gr.ix.ntotal, 0), T(gr.ix.M * 256) {}

    void set_query(const float *x)
    {
        const HostIndex &ix = g.ix;
        for (uint64_t m = 0; m < ix.M; ++m)
            for (int j = 0; j < 256; ++j) {
                const float *c = &ix.centroids[(m * 256 + j) * ix.dsub];
                float acc = 0;
                for (uint64_t t = 0; t < ix.dsub; ++t) { float e = x[m * ix.dsub + t] - c[t]; acc = fmaf(e, e, acc); }
                T[m * 256 + j] = acc;
            }
    }
    float dis(int64_t v) const
    {
        const uint8_t *c = &g.ix.codes[(size_t)v * g.ix.code_size];
        float s = 0;
        for (uint64_t m = 0; m < g.ix.M; ++m) s += T[m * 256 + c[m]];
        return s;
    }
    void next_stamp()
    {
        if (++stamp == 0) { std::fill(visited.begin(), visited.end(), 0); stamp = 1; }
    }

    // add_link(src -> dest) under src's lock
    void add_link(int64_t src, int32_t dest, int level)
    {
        int width;
        int32_t *r = g.row(src, level, width);
        for (int j = 0; j < width; ++j) {
            if (r[j] == dest) return;
            if (r[j] < 0) { r[j] = dest; return; }
        }
        tmp.clear();
        tmp.push_back({g.sym(src, dest), dest});
        for (int j = 0; j < width; ++j) tmp.push_back({g.sym(src, r[j]), r[j]});
        std::sort(tmp.begin(), tmp.end(), [](const Cand &a, const Cand &b) { return a.d < b.d || (a.d == b.d && a.id < b.id); });
        shrink(g, tmp, (size_t)width, sel);
        for (int j = 0; j < width; ++j) r[j] = j < (int)sel.size() ? sel[(size_t)j].id : -1;
    }

    void insert(int64_t pt, int pt_level, const float *x, int32_t &entry_point, int &max_level, int efc)
    {
        set_query(x);
        int32_t nearest = entry_point;
        if (nearest < 0) return;
        float d_nearest = dis(nearest);
        int level = max_level;
        for (; level > pt_level; --level) {
            for (;;) {
                int32_t prev = nearest;
                int width;
                const int32_t *r = g.row(nearest, level, width);
                for (int j = 0; j < width; ++j) {
                    int32_t v = r[j];
                    if (v < 0) break;
                    float d = dis(v);
                    if (d < d_nearest) { nearest = v; d_nearest = d; }
                }
                if (nearest == prev) break;
            }
        }
        for (; level >= 0; --level) {
            // search_neighbors_to_add
            next_stamp();
            std::priority_queue<Cand, std::vector<Cand>, Farther> results;
            std::priority_queue<Cand, std::vector<Cand>, Closer> cands;
            results.push({d_nearest, nearest});
            cands.push({d_nearest, nearest});
            visited[(size_t)nearest] = stamp;
            visited[(size_t)pt] = stamp;
            while (!cands.empty()) {
                Cand c = cands.top();
                if (c.d > results.top().d) break;
                cands.pop();
                int width;
                const int32_t *r = g.row(c.id, level, width);
                for (int j = 0; j < width; ++j) {
                    int32_t v = r[j];
                    if (v < 0) break;
                    if (visited[(size_t)v] == stamp) continue;
                    visited[(size_t)v] = stamp;
                    float d = dis(v);
                    if ((int)results.size() < efc || d < results.top().d) {
                        results.push({d, v});
                        cands.push({d, v});
                        if ((int)results.size() > efc) results.pop();
                    }
                }
            }
            std::vector<Cand> sorted(results.size());
            for (size_t i = results.size(); i-- > 0;) { sorted[i] = results.top(); results.pop(); }
            int width;
            g.row(pt, level, width);
            std::vector<Cand> chosen;
            shrink(g, sorted, (size_t)width, chosen);
            omp_set_lock(&g.locks[(size_t)pt]);
            for (const Cand &c : chosen) add_link(pt, c.id, level);
            omp_unset_lock(&g.locks[(size_t)pt]);
            for (const Cand &c : chosen) {
                omp_set_lock(&g.locks[(size_t)c.id]);
                add_link(c.id, (int32_t)pt, level);
                omp_unset_lock(&g.locks[(size_t)c.id]);
            }
            if (!sorted.empty()) { nearest = sorted[0].id; d_nearest = sorted[0].d; }
        }
    }
};

}  // namespace

// HNSW level law (HNSW::set_default_probas), random levels, offsets and the empty neighbour table.
void builder_level_law(HostIndex &ix, int64_t n, int M_hnsw, uint32_t seed)
{
    const double level_mult = 1.0 / std::log((double)M_hnsw);
    ix.assign_probas.clear();
    ix.cum.assign(1, 0);
    int nn = 0;
    for (int level = 0;; ++level) {
        double proba = std::exp(-level / level_mult) * (1 - std::exp(-1 / level_mult));
        if (proba < 1e-9) break;
        ix.assign_probas.push_back(proba);
        nn += level == 0 ? 2 * M_hnsw : M_hnsw;
        ix.cum.push_back(nn);
    }
    ix.levels.resize((size_t)n);
    {
        std::mt19937 rng(seed ^ 0x9e3779b9u);
        std::uniform_real_distribution<double> uni(0.0, 1.0);
        for (int64_t i = 0; i < n; ++i) {
            double f = uni(rng);
            int lv = (int)ix.assign_probas.size() - 1;
            for (size_t l = 0; l < ix.assign_probas.size(); ++l) {
                if (f < ix.assign_probas[l]) { lv = (int)l; break; }
                f -= ix.assign_probas[l];
            }
            ix.levels[(size_t)i] = lv + 1;
        }
    }
    ix.offsets.resize((size_t)n + 1);
    ix.offsets[0] = 0;
    for (int64_t i = 0; i < n; ++i) ix.offsets[(size_t)i + 1] = ix.offsets[(size_t)i] + (uint64_t)ix.cum[(size_t)ix.levels[(size_t)i]];
    ix.neighbors.assign((size_t)ix.offsets[(size_t)n], -1);
}

// Insertion order: descending level, random inside a level (hnsw_add_vertices).
void builder_insertion_order(const HostIndex &ix, uint32_t seed, std::vector<int64_t> &order)
{
    order.resize((size_t)ix.ntotal);
    std::iota(order.begin(), order.end(), 0);
    std::mt19937 rng(seed + 77);
    std::shuffle(order.begin(), order.end(), rng);
    std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return ix.levels[(size_t)a] > ix.levels[(size_t)b]; });
}

// Rows of the PQ training set: at most 256 points per centroid, a seeded shuffle of the sample when it is larger.
void builder_training_pick(int64_t n_train, uint32_t seed, std::vector<int64_t> &pick)
{
    const int64_t nt = std::min<int64_t>(n_train, 256 * 256);
    pick.resize((size_t)n_train);
    std::iota(pick.begin(), pick.end(), 0);
    std::mt19937 rng(seed);
    if (nt < n_train) std::shuffle(pick.begin(), pick.end(), rng);
    pick.resize((size_t)nt);
}

// k-means start: 256 seeded random rows of the training set (row c % n for tiny sets).
void builder_kmeans_init_rows(int64_t n, int k, uint32_t seed, std::vector<int64_t> &rows)
{
    std::mt19937 rng(seed);
    std::vector<int64_t> perm((size_t)n);
    std::iota(perm.begin(), perm.end(), 0);
    std::shuffle(perm.begin(), perm.end(), rng);
    rows.resize((size_t)k);
    for (int c = 0; c < k; ++c) rows[(size_t)c] = perm[(size_t)(c % n)];
}

// Centroid update of one k-means iteration from per-cluster sums and counts, with FAISS's empty-cluster split.
void builder_kmeans_update(const double *sum, int64_t *cnt, int k, int dsub, float *cent)
{
    for (int c = 0; c < k; ++c)
        if (cnt[c])
            for (int t = 0; t < dsub; ++t) cent[(size_t)c * dsub + t] = (float)(sum[(size_t)c * dsub + t] / (double)cnt[c]);
    for (int c = 0; c < k; ++c) {
        if (cnt[c]) continue;
        int big = (int)(std::max_element(cnt, cnt + k) - cnt);
        for (int t = 0; t < dsub; ++t) {
            float v = cent[(size_t)big * dsub + t];
            float eps = 1.f / 1024.f;
            cent[(size_t)c * dsub + t] = v * (1 + ((t & 1) ? eps : -eps));
            cent[(size_t)big * dsub + t] = v * (1 - ((t & 1) ? eps : -eps));
        }
        cnt[c] = cnt[big] / 2;
        cnt[big] -= cnt[c];
    }
}

// x: n x d row-major.  train: n_train x d (the reference passes an evenly strided 50 % sample).
int build_index_cpu(const float *x, int64_t n, int d, const float *train, int64_t n_train, int M_pq, int nbits, int M_hnsw,
                    int efc, int threads, uint32_t seed, HostIndex &ix)
{
    if (n <= 0 || d <= 0 || M_pq <= 0 || d % M_pq) return fail(DRM_EINVAL, "bad index build arguments (d must be a multiple of M_pq)");
    if (nbits != 8) return fail(DRM_EUNSUPPORTED, "only nbits == 8 is supported by the builder");
    if (M_hnsw < 2 || efc < 1) return fail(DRM_EINVAL, "M_hnsw must be >= 2 and EFC >= 1");
    if (n >= ((int64_t)1 << 31) - 1) return fail(DRM_EUNSUPPORTED, "more than 2^31 vectors");
    if (threads <= 0) threads = omp_get_max_threads();
    omp_set_num_threads(threads);
    ix = HostIndex();
    ix.d = d; ix.ntotal = n; ix.M = (uint64_t)M_pq; ix.nbits = 8; ix.ksub = 256; ix.dsub = (uint64_t)(d / M_pq); ix.code_size = (uint64_t)M_pq;
    ix.efConstruction = efc; ix.efSearch = 16; ix.upper_beam = 1;
    const int dsub = d / M_pq;

    // ---- ProductQuantizer::train: k-means per sub-space on at most 256 points per centroid
    ix.centroids.assign((size_t)M_pq * 256 * dsub, 0.f);
    {
        if (!train || n_train <= 0) { train = x; n_train = n; }
        std::vector<int64_t> pick;
        builder_training_pick(n_train, seed, pick);
        const int64_t nt = (int64_t)pick.size();
        std::vector<float> sub((size_t)nt * dsub);
        for (int m = 0; m < M_pq; ++m) {
            for (int64_t i = 0; i < nt; ++i)
                memcpy(&sub[(size_t)i * dsub], train + (size_t)pick[(size_t)i] * d + (size_t)m * dsub, sizeof(float) * dsub);
            kmeans(sub.data(), nt, dsub, 256, 25, seed + 1 + (uint32_t)m, &ix.centroids[(size_t)m * 256 * dsub]);
        }
    }
    // ---- IndexPQ::add: nearest centroid per sub-space
    ix.codes.resize((size_t)n * M_pq);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i)
        for (int m = 0; m < M_pq; ++m) {
            const float *xi = x + (size_t)i * d + (size_t)m * dsub;
            float best = FLT_MAX; int bj = 0;
            for (int j = 0; j < 256; ++j) {
                const float *c = &ix.centroids[((size_t)m * 256 + j) * dsub];
                float s = 0;
                for (int t = 0; t < dsub; ++t) { float e = xi[t] - c[t]; s += e * e; }
                if (s < best) { best = s; bj = j; }
            }
            ix.codes[(size_t)i * M_pq + m] = (uint8_t)bj;
        }

    builder_level_law(ix, n, M_hnsw, seed);

    Graph g(ix);
    g.sdc.resize((size_t)M_pq * 256 * 256);
#pragma omp parallel for schedule(static)
    for (int mj = 0; mj < M_pq * 256; ++mj) {
        int m = mj / 256, j = mj % 256;
        const float *a = &ix.centroids[((size_t)m * 256 + j) * dsub];
        for (int j2 = 0; j2 < 256; ++j2) {
            const float *b = &ix.centroids[((size_t)m * 256 + j2) * dsub];
            float s = 0;
            for (int t = 0; t < dsub; ++t) { float e = a[t] - b[t]; s += e * e; }
            g.sdc[((size_t)m * 256 + j) * 256 + j2] = s;
        }
    }
    g.locks.resize((size_t)n);
    for (auto &l : g.locks) omp_init_lock(&l);

    // ---- insertion order: descending level, random inside a level (hnsw_add_vertices)
    std::vector<int64_t> order;
    builder_insertion_order(ix, seed, order);
    int32_t entry_point = (int32_t)order[0];
    int max_level = ix.levels[(size_t)order[0]] - 1;
    size_t begin = 1;
    while (begin < (size_t)n) {
        int lv = ix.levels[(size_t)order[begin]];
        size_t end = begin;
        while (end < (size_t)n && ix.levels[(size_t)order[end]] == lv) ++end;
#pragma omp parallel
        {
            Worker w(g);
#pragma omp for schedule(dynamic, 64)
            for (int64_t i = (int64_t)begin; i < (int64_t)end; ++i) {
                int64_t pt = order[(size_t)i];
                int32_t ep = entry_point; int ml = max_level;  // fixed while a level group is inserted
                w.insert(pt, lv - 1, x + (size_t)pt * d, ep, ml, efc);
            }
        }
        begin = end;
    }
    for (auto &l : g.locks) omp_destroy_lock(&l);
    ix.entry_point = entry_point;
    ix.max_level = max_level;
    return DRM_OK;
}

}  // namespace drm

extern "C" int drm_build_index(const float *x, int64_t n, int d, const float *train, int64_t n_train, int M_pq, int nbits,
                               int M_hnsw, int efc, int threads, uint32_t seed, const char *out_path)
{
    if (!x || !out_path) return drm::fail(DRM_EINVAL, "null argument to drm_build_index");
    drm::HostIndex ix;
    int rc = drm::build_index_cpu(x, n, d, train, n_train, M_pq, nbits, M_hnsw, efc, threads, seed, ix);
    if (rc != DRM_OK) return rc;
    return drm::write_index_file(out_path, ix);
}
