/*
 * TEST INFRASTRUCTURE ONLY — the CPU oracle for the HNSW-PQ search half of the hot path.
 * Nothing in the product links, loads or calls this file; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference leg do.
 *
 * PARITY UNPINNED.  The arithmetic of this half lives in FAISS (faiss::IndexHNSWPQ), a third-party
 * dependency that is NOT under /root/reference: conda package `faiss-cpu`, channel conda-forge, no
 * version pin (environment.yml:2-13), linked as -lfaiss (build.zig:87, CMakeLists.txt:29-47).  It is
 * not installed in this image and the reference ships no index file, no expected indices/distances
 * and no recall figure, so nothing here can be checked against the real library.
 * This file restates FAISS's published algorithm (faiss >= 1.8: faiss/impl/HNSW.cpp `HNSW::search`,
 * `greedy_update_nearest`, `search_from_candidates`, `MinimaxHeap`; faiss/utils/Heap.h `heap_push`,
 * `heap_pop`, `heap_replace_top`, `heap_reorder` with `CMax::cmp2`; faiss/IndexPQ.cpp
 * `PQDistanceComputer`; faiss/impl/index_write.cpp `IHNp`/`IxPq`) and is anchored on the reference's
 * call sites:
 *   faiss::read_index + dynamic_cast<IndexHNSWPQ*>   src/main.cpp:236-237, src/hnswpq/test_search.cpp:68-76
 *   index->hnsw.efSearch = ef; index->search(n,x,k,D,I)   src/hnswpq/search.cpp:13,39-40
 *   IndexHNSWPQ(dim, M_pq, M_hnsw, nbits), write_index  src/hnswpq/index.cpp:111-113,188
 *
 * One deliberate knob: `tie_rule`.
 *   0 = FAISS    MinimaxHeap::pop_min breaks exact distance ties by the highest slot of FAISS's
 *                binary-heap array (layout dependent).
 *   1 = CANONICAL  every distance comparison is made on the total order (distance, id) — see
 *                search_one_canonical below.  This is the rule the CUDA kernel implements (a
 *                layout-free sorted pool), so GPU == oracle(1) bit for bit, and oracle(0) vs
 *                oracle(1) measures what the tie rule changes (only queries where two candidates
 *                have bit-equal PQ distances — BASELINE.json's "identical except where PQ
 *                distances tie").
 * The ADC table uses a fixed summation order with fmaf so that CPU and GPU agree to the last bit
 * (real FAISS uses SIMD reductions whose last ulp differs).
 */
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <atomic>
#include <cstring>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

thread_local std::string g_err;

struct OrcIndex {
    int32_t d = 0;
    int64_t ntotal = 0;
    int32_t metric = 1;
    // HNSW
    std::vector<double> assign_probas;
    std::vector<int32_t> cum;  // cum_nneighbor_per_level
    std::vector<int32_t> levels;
    std::vector<uint64_t> offsets;
    std::vector<int32_t> neighbors;
    int32_t entry_point = -1, max_level = -1, efConstruction = 0, efSearch = 0;
    // PQ
    uint64_t pq_d = 0, M = 0, nbits = 0, ksub = 0, dsub = 0, code_size = 0;
    std::vector<float> centroids;  // [M][ksub][dsub]
    std::vector<uint8_t> codes;    // ntotal * code_size
};

struct Reader {
    FILE *f;
    bool ok = true;
    template <class T> T one() {
        T v{};
        if (fread(&v, sizeof(T), 1, f) != 1) ok = false;
        return v;
    }
    template <class T> bool vec(std::vector<T> &v, uint64_t max_elems) {
        uint64_t n = one<uint64_t>();
        if (!ok || n > max_elems) { ok = false; return false; }
        v.resize((size_t)n);
        if (n && fread(v.data(), sizeof(T), (size_t)n, f) != n) ok = false;
        return ok;
    }
};

uint32_t fourcc(const char *s) { return (uint32_t)(uint8_t)s[0] | (uint32_t)(uint8_t)s[1] << 8 | (uint32_t)(uint8_t)s[2] << 16 | (uint32_t)(uint8_t)s[3] << 24; }

bool read_header(Reader &r, int32_t &d, int64_t &ntotal, int32_t &metric)
{
    d = r.one<int32_t>();
    ntotal = r.one<int64_t>();
    r.one<int64_t>();
    r.one<int64_t>();
    r.one<uint8_t>();  // is_trained
    metric = r.one<int32_t>();
    if (metric > 1) r.one<float>();
    return r.ok;
}

// FAISS PQDecoderGeneric: codes are a little-endian bit string, nbits per sub-quantizer.
inline uint32_t decode_code(const uint8_t *code, uint64_t nbits, uint64_t m)
{
    if (nbits == 8) return code[m];
    uint64_t bit = m * nbits;
    uint32_t v = 0;
    for (uint64_t b = 0; b < nbits; ++b, ++bit) v |= (uint32_t)((code[bit >> 3] >> (bit & 7)) & 1u) << b;
    return v;
}

// T[m][j] = sum_t (x[m*dsub+t] - c[m][j][t])^2, t ascending, one fmaf per term.
void adc_table(const OrcIndex &ix, const float *x, float *T)
{
    for (uint64_t m = 0; m < ix.M; ++m)
        for (uint64_t j = 0; j < ix.ksub; ++j) {
            const float *c = &ix.centroids[(m * ix.ksub + j) * ix.dsub];
            float acc = 0.f;
            for (uint64_t t = 0; t < ix.dsub; ++t) {
                float diff = x[m * ix.dsub + t] - c[t];
                acc = fmaf(diff, diff, acc);
            }
            T[m * ix.ksub + j] = acc;
        }
}

inline float adc_dis(const OrcIndex &ix, const float *T, int64_t v)
{
    const uint8_t *code = &ix.codes[(uint64_t)v * ix.code_size];
    float acc = 0.f;
    for (uint64_t m = 0; m < ix.M; ++m) acc += T[m * ix.ksub + decode_code(code, ix.nbits, m)];
    return acc;
}

// ---- faiss/utils/Heap.h, CMax<float, idx> with cmp2 (ties broken on id) ----
template <class TI> inline bool cmp2(float a1, float b1, TI a2, TI b2) { return (a1 > b1) || ((a1 == b1) && (a2 > b2)); }

template <class TI> void heap_pop(size_t k, float *val, TI *ids)
{
    val--; ids--;
    float v = val[k];
    TI id = ids[k];
    size_t i = 1, i1, i2;
    for (;;) {
        i1 = i << 1;
        i2 = i1 + 1;
        if (i1 > k) break;
        if (i2 == k + 1 || cmp2(val[i1], val[i2], ids[i1], ids[i2])) {
            if (cmp2(v, val[i1], id, ids[i1])) break;
            val[i] = val[i1]; ids[i] = ids[i1]; i = i1;
        } else {
            if (cmp2(v, val[i2], id, ids[i2])) break;
            val[i] = val[i2]; ids[i] = ids[i2]; i = i2;
        }
    }
    val[i] = val[k]; ids[i] = ids[k];
}

template <class TI> void heap_push(size_t k, float *val, TI *ids, float v, TI id)
{
    val--; ids--;
    size_t i = k, f;
    while (i > 1) {
        f = i >> 1;
        if (!cmp2(v, val[f], id, ids[f])) break;
        val[i] = val[f]; ids[i] = ids[f]; i = f;
    }
    val[i] = v; ids[i] = id;
}

template <class TI> void heap_replace_top(size_t k, float *val, TI *ids, float v, TI id)
{
    val--; ids--;
    size_t i = 1, i1, i2;
    for (;;) {
        i1 = i << 1;
        i2 = i1 + 1;
        if (i1 > k) break;
        if (i2 == k + 1 || cmp2(val[i1], val[i2], ids[i1], ids[i2])) {
            if (cmp2(v, val[i1], id, ids[i1])) break;
            val[i] = val[i1]; ids[i] = ids[i1]; i = i1;
        } else {
            if (cmp2(v, val[i2], id, ids[i2])) break;
            val[i] = val[i2]; ids[i] = ids[i2]; i = i2;
        }
    }
    val[i] = v; ids[i] = id;
}

template <class TI> size_t heap_reorder(size_t k, float *val, TI *ids)
{
    size_t i, ii;
    for (i = 0, ii = 0; i < k; i++) {
        float v = val[0];
        TI id = ids[0];
        heap_pop<TI>(k - i, val, ids);
        val[k - ii - 1] = v;
        ids[k - ii - 1] = id;
        if (id != -1) ii++;
    }
    size_t nel = ii;
    memmove(val, val + k - ii, ii * sizeof(*val));
    memmove(ids, ids + k - ii, ii * sizeof(*ids));
    for (; ii < k; ii++) { val[ii] = FLT_MAX; ids[ii] = -1; }
    return nel;
}

// analysis counters (tools only), see orc_tie_counters
static std::atomic<long long> g_tie_queries{0}, g_tie_total{0}, g_tie_popped_queries{0};

// ---- HNSW::MinimaxHeap ----
struct MinimaxHeap {
    int n, k = 0, nvalid = 0;
    std::vector<int32_t> ids;
    std::vector<float> dis;
    explicit MinimaxHeap(int n_) : n(n_), ids((size_t)n_), dis((size_t)n_) {}
    int tie_evict = 0, tie_evict_popped = 0;  // analysis: evictions whose worst distance is shared / shared with a popped entry
    void push(int32_t i, float v)
    {
        if (k == n) {
            if (v >= dis[0]) return;
            {
                bool tie = false, popped = ids[0] == -1;
                for (int c = 1; c <= 2 && c < k; ++c)
                    if (dis[(size_t)c] == dis[0]) { tie = true; popped = popped || ids[(size_t)c] == -1; }
                if (tie) { tie_evict++; if (popped) { tie_evict_popped++; } }
            }
            if (ids[0] != -1) --nvalid;
            heap_pop<int32_t>((size_t)k--, dis.data(), ids.data());
        }
        heap_push<int32_t>((size_t)++k, dis.data(), ids.data(), v, i);
        ++nvalid;
    }
    int size() const { return nvalid; }
    int pop_min(float *vmin_out, int tie_rule)
    {
        int imin = -1;
        float vmin = 0.f;
        if (tie_rule == 0) {
            // FAISS: scan from the last slot down, strict '<' => highest slot among equal minima.
            int i = k - 1;
            while (i >= 0 && ids[(size_t)i] == -1) i--;
            if (i == -1) return -1;
            imin = i; vmin = dis[(size_t)i];
            for (i--; i >= 0; i--)
                if (ids[(size_t)i] != -1 && dis[(size_t)i] < vmin) { vmin = dis[(size_t)i]; imin = i; }
        } else {
            // CANONICAL: smallest (dis, id) among live entries — independent of the array layout.
            for (int i = 0; i < k; ++i) {
                if (ids[(size_t)i] == -1) continue;
                if (imin < 0 || dis[(size_t)i] < vmin || (dis[(size_t)i] == vmin && ids[(size_t)i] < ids[(size_t)imin])) {
                    imin = i; vmin = dis[(size_t)i];
                }
            }
            if (imin < 0) return -1;
        }
        if (vmin_out) *vmin_out = vmin;
        int ret = ids[(size_t)imin];
        ids[(size_t)imin] = -1;
        --nvalid;
        return ret;
    }
    int count_below(float thresh) const
    {
        int c = 0;
        for (int i = 0; i < k; ++i) c += dis[(size_t)i] < thresh;
        return c;
    }
};

struct VisitedTable {
    std::vector<uint8_t> visited;
    uint8_t visno = 1;
    explicit VisitedTable(size_t n) : visited(n, 0) {}
    void set(int64_t v) { visited[(size_t)v] = visno; }
    bool get(int64_t v) const { return visited[(size_t)v] == visno; }
    void advance()
    {
        visno++;
        if (visno == 250) { std::fill(visited.begin(), visited.end(), 0); visno = 1; }
    }
};

struct Stats { int64_t ndis = 0, nhops = 0, ndis_up = 0, nhops_up = 0; };

// One query of IndexHNSW::search -> HNSW::search (bounded queue, check_relative_distance, no selector).
void search_one(const OrcIndex &ix, const float *x, int k, int efSearch, int tie_rule, VisitedTable &vt,
                std::vector<float> &T, float *D, int64_t *I, Stats &st)
{
    // result max-heap initialised to (FLT_MAX, -1)
    for (int j = 0; j < k; ++j) { D[j] = FLT_MAX; I[j] = -1; }
    if (ix.entry_point < 0 || ix.ntotal == 0) return;
    adc_table(ix, x, T.data());

    // greedy descent on levels max_level..1 (greedy_update_nearest)
    int32_t nearest = ix.entry_point;
    float d_nearest = adc_dis(ix, T.data(), nearest);
    for (int level = ix.max_level; level >= 1; --level) {
        for (;;) {
            int32_t prev = nearest;
            uint64_t o = ix.offsets[(size_t)nearest];
            uint64_t begin = o + (uint64_t)ix.cum[(size_t)level], end = o + (uint64_t)ix.cum[(size_t)level + 1];
            for (uint64_t j = begin; j < end; ++j) {
                int32_t v = ix.neighbors[j];
                if (v < 0) break;
                float dis = adc_dis(ix, T.data(), v);
                st.ndis_up++;
                if (dis < d_nearest) { nearest = v; d_nearest = dis; }
            }
            st.nhops_up++;
            if (nearest == prev) break;
        }
    }

    // level 0: search_from_candidates
    int ef = std::max(efSearch, k);
    MinimaxHeap cand(ef);
    cand.push(nearest, d_nearest);
    float threshold = D[0];
    for (int i = 0; i < cand.k; ++i) {
        int32_t v1 = cand.ids[(size_t)i];
        float dd = cand.dis[(size_t)i];
        if (dd < threshold) { heap_replace_top<int64_t>((size_t)k, D, I, dd, v1); threshold = D[0]; }
        vt.set(v1);
    }
    while (cand.size() > 0) {
        float d0 = 0;
        int v0 = cand.pop_min(&d0, tie_rule);
        if (cand.count_below(d0) >= efSearch) break;
        uint64_t o = ix.offsets[(size_t)v0];
        uint64_t begin = o + (uint64_t)ix.cum[0], end = o + (uint64_t)ix.cum[1];
        for (uint64_t j = begin; j < end; ++j) {
            int32_t v1 = ix.neighbors[j];
            if (v1 < 0) break;
            if (vt.get(v1)) continue;
            vt.set(v1);
            float dd = adc_dis(ix, T.data(), v1);
            st.ndis++;
            if (dd < threshold) { heap_replace_top<int64_t>((size_t)k, D, I, dd, v1); threshold = D[0]; }
            cand.push(v1, dd);
        }
        st.nhops++;
    }
    vt.advance();
    heap_reorder<int64_t>((size_t)k, D, I);
    g_tie_total++;
    if (cand.tie_evict) g_tie_queries++;
    if (cand.tie_evict_popped) g_tie_popped_queries++;
}

// CANONICAL tie rule (tie_rule = 1): the same search, but every comparison that FAISS makes on the
// distance alone is made on the total order (distance, id), and a popped candidate keeps its id:
//   push      : a full candidate set rejects v unless (d, v) < its largest (d, id) entry;
//   pop_min   : the unexpanded entry with the smallest (d, id);
//   count_below(d0): entries with d < d0 (unchanged);
//   results   : the k smallest (d, id) among all evaluated vertices.
// Written out, the candidate MinimaxHeap becomes a sorted array of at most cap = max(ef, k) entries
// and - because cap >= k - the result heap is simply its first k entries.  Wherever no two distances
// are bit-equal this is exactly FAISS's procedure; it differs from tie_rule 0 only in which of
// several equal-distance vertices survive at a full boundary.  The CUDA kernel implements this rule.
// Analysis counters (tools only): how often the SECOND unexpanded pool entry at the start of a hop is the entry the next
// hop expands - i.e. how often a kernel that evaluated two rows at once could keep both results without changing the
// search (orc_spec2_counters).
static std::atomic<long long> g_spec2_yes{0}, g_spec2_total{0};
// Analysis counter (tools only): queries of the canonical search in which at least one decision compared two EQUAL
// distances (pop order among equal minima, a candidate equal to the worst kept distance, an eviction among equal worst
// distances) - the only places where FAISS's rule can decide differently (orc_tie_counters).


void search_one_canonical(const OrcIndex &ix, const float *x, int k, int efSearch, VisitedTable &vt,
                          std::vector<float> &T, float *D, int64_t *I, Stats &st)
{
    for (int j = 0; j < k; ++j) { D[j] = FLT_MAX; I[j] = -1; }
    if (ix.entry_point < 0 || ix.ntotal == 0) return;
    adc_table(ix, x, T.data());
    int32_t nearest = ix.entry_point;
    float d_nearest = adc_dis(ix, T.data(), nearest);
    for (int level = ix.max_level; level >= 1; --level) {
        for (;;) {
            int32_t prev = nearest;
            uint64_t o = ix.offsets[(size_t)nearest];
            uint64_t begin = o + (uint64_t)ix.cum[(size_t)level], end = o + (uint64_t)ix.cum[(size_t)level + 1];
            for (uint64_t j = begin; j < end; ++j) {
                int32_t v = ix.neighbors[j];
                if (v < 0) break;
                float dis = adc_dis(ix, T.data(), v);
                st.ndis_up++;
                if (dis < d_nearest) { nearest = v; d_nearest = dis; }
            }
            st.nhops_up++;
            if (nearest == prev) break;
        }
    }
    struct Ent { float d; int32_t id; bool expanded; };
    auto less = [](float d1, int32_t i1, float d2, int32_t i2) { return d1 < d2 || (d1 == d2 && i1 < i2); };
    const size_t cap = (size_t)std::max(efSearch, k);
    std::vector<Ent> pool;
    pool.reserve(cap + 1);
    pool.push_back({d_nearest, nearest, false});
    vt.set(nearest);
    bool tie_event = false;
    for (;;) {
        size_t cur = 0;
        while (cur < pool.size() && pool[cur].expanded) ++cur;
        if (cur == pool.size()) break;
        const float d0 = pool[cur].d;
        {   // tie analysis (a): another unexpanded entry at the same distance
            size_t nx = cur + 1;
            while (nx < pool.size() && pool[nx].expanded) ++nx;
            if (nx < pool.size() && pool[nx].d == d0) tie_event = true;
        }
        const int32_t v0 = pool[cur].id;
        pool[cur].expanded = true;
        {   // spec-2 analysis: was this hop's vertex the runner-up of the previous hop?
            static thread_local int32_t runner_up = -1;
            if (st.nhops > 0 || runner_up >= 0) { g_spec2_total++; if (runner_up == v0) g_spec2_yes++; }
            size_t nx = cur + 1;
            while (nx < pool.size() && pool[nx].expanded) ++nx;
            runner_up = nx < pool.size() ? pool[nx].id : -1;
            if (cur == 0 && pool.size() == 1) runner_up = -1;
        }
        int below = 0;
        for (const Ent &e : pool) below += e.d < d0;
        if (below >= efSearch) break;
        uint64_t o = ix.offsets[(size_t)v0];
        uint64_t begin = o + (uint64_t)ix.cum[0], end = o + (uint64_t)ix.cum[1];
        for (uint64_t j = begin; j < end; ++j) {
            int32_t v1 = ix.neighbors[j];
            if (v1 < 0) break;
            if (vt.get(v1)) continue;
            vt.set(v1);
            float dd = adc_dis(ix, T.data(), v1);
            st.ndis++;
            if (pool.size() == cap && dd == pool.back().d) tie_event = true;                       // (b) equal to the worst kept
            if (pool.size() >= (size_t)k && dd == pool[(size_t)k - 1].d) tie_event = true;          // (c) equal to the k-th result
            if (pool.size() == cap && !less(dd, v1, pool.back().d, pool.back().id)) continue;
            size_t pos = 0;
            while (pos < pool.size() && less(pool[pos].d, pool[pos].id, dd, v1)) ++pos;
            pool.insert(pool.begin() + (ptrdiff_t)pos, Ent{dd, v1, false});
            if (pool.size() > cap) {
                if (pool[cap].d == pool[cap - 1].d) tie_event = true;                               // (b) eviction among equal worst
                pool.pop_back();
            }
            if (pool.size() > (size_t)k && pool[(size_t)k].d == pool[(size_t)k - 1].d) tie_event = true;  // (c) result cut inside a tie class
        }
        st.nhops++;
    }
    vt.advance();
    g_tie_total++;
    if (tie_event) g_tie_queries++;
    for (size_t j = 0; j < pool.size() && j < (size_t)k; ++j) { D[j] = pool[j].d; I[j] = pool[j].id; }
}

}  // namespace

extern "C" {

const char *orc_last_error() { return g_err.c_str(); }
void orc_spec2_counters(int64_t *out, int reset)
{
    out[0] = g_spec2_yes.load(); out[1] = g_spec2_total.load();
    if (reset) { g_spec2_yes = 0; g_spec2_total = 0; }
}

void orc_tie_counters(int64_t *out, int reset)
{
    out[0] = g_tie_queries.load(); out[1] = g_tie_total.load(); out[2] = g_tie_popped_queries.load();
    if (reset) { g_tie_queries = 0; g_tie_total = 0; g_tie_popped_queries = 0; }
}

void *orc_index_read(const char *path)
{
    FILE *f = fopen(path, "rb");
    if (!f) { g_err = std::string("cannot open ") + path; return nullptr; }
    Reader r{f};
    auto *ix = new OrcIndex();
    auto fail = [&](const char *why) { g_err = why; fclose(f); delete ix; return (void *)nullptr; };
    if (r.one<uint32_t>() != fourcc("IHNp")) return fail("not an IHNp (IndexHNSWPQ) file");
    if (!read_header(r, ix->d, ix->ntotal, ix->metric)) return fail("truncated header");
    if (ix->d <= 0 || ix->ntotal < 0) return fail("bad header");
    const uint64_t big = (uint64_t)1 << 40;
    if (!r.vec(ix->assign_probas, 1024) || !r.vec(ix->cum, 1024) || !r.vec(ix->levels, big) ||
        !r.vec(ix->offsets, big) || !r.vec(ix->neighbors, big))
        return fail("truncated HNSW block");
    ix->entry_point = r.one<int32_t>();
    ix->max_level = r.one<int32_t>();
    ix->efConstruction = r.one<int32_t>();
    ix->efSearch = r.one<int32_t>();
    r.one<int32_t>();  // upper_beam
    if (r.one<uint32_t>() != fourcc("IxPq")) return fail("storage is not IxPq (IndexPQ)");
    int32_t d2, metric2; int64_t n2;
    if (!read_header(r, d2, n2, metric2)) return fail("truncated storage header");
    ix->pq_d = r.one<uint64_t>(); ix->M = r.one<uint64_t>(); ix->nbits = r.one<uint64_t>();
    if (!r.ok || ix->M == 0 || ix->nbits == 0 || ix->nbits > 16 || ix->pq_d != (uint64_t)ix->d || ix->pq_d % ix->M)
        return fail("bad ProductQuantizer header");
    ix->ksub = (uint64_t)1 << ix->nbits; ix->dsub = ix->pq_d / ix->M; ix->code_size = (ix->nbits * ix->M + 7) / 8;
    if (!r.vec(ix->centroids, big) || !r.vec(ix->codes, big)) return fail("truncated PQ block");
    r.one<int32_t>(); r.one<uint8_t>(); r.one<int32_t>();  // search_type, encode_signs, polysemous_ht
    fclose(f);
    if (!r.ok) { g_err = "truncated tail"; delete ix; return nullptr; }
    if ((int64_t)ix->levels.size() != ix->ntotal || (int64_t)ix->offsets.size() != ix->ntotal + 1 ||
        ix->centroids.size() != ix->M * ix->ksub * ix->dsub || ix->codes.size() != (uint64_t)ix->ntotal * ix->code_size ||
        n2 != ix->ntotal || d2 != ix->d) {
        g_err = "inconsistent array lengths"; delete ix; return nullptr;
    }
    return ix;
}

void orc_index_free(void *h) { delete (OrcIndex *)h; }

// info[0..9] = d, ntotal, M, nbits, M_hnsw (cum[1]/2), entry_point, max_level, efConstruction, efSearch, neighbors.size()
void orc_index_info(void *h, int64_t *info)
{
    auto *ix = (OrcIndex *)h;
    info[0] = ix->d; info[1] = ix->ntotal; info[2] = (int64_t)ix->M; info[3] = (int64_t)ix->nbits;
    info[4] = ix->cum.size() > 1 ? ix->cum[1] / 2 : 0; info[5] = ix->entry_point; info[6] = ix->max_level;
    info[7] = ix->efConstruction; info[8] = ix->efSearch; info[9] = (int64_t)ix->neighbors.size();
}

// stats[0..3] = sum ndis (level 0), sum nhops (level 0), sum ndis (upper), sum nhops (upper);
// per_query (nullable) gets n x 2 int32 (ndis, nhops) for the byte model.
int orc_search(void *h, const float *x, int64_t n, int k, int ef, int tie_rule, float *D, int64_t *I,
               int64_t *stats, int32_t *per_query, int threads)
{
    auto *ix = (OrcIndex *)h;
    if (n <= 0) { g_err = "Query data is empty"; return -1; }
    if (threads <= 0) {
#ifdef _OPENMP
        threads = omp_get_max_threads();
#else
        threads = 1;
#endif
    }
    int64_t s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma omp parallel num_threads(threads) reduction(+ : s0, s1, s2, s3)
    {
        VisitedTable vt((size_t)ix->ntotal);
        std::vector<float> T(ix->M * ix->ksub);
#pragma omp for schedule(guided)
        for (int64_t i = 0; i < n; ++i) {
            Stats st;
            if (tie_rule == 1) search_one_canonical(*ix, x + i * ix->d, k, ef, vt, T, D + i * k, I + i * k, st);
            else search_one(*ix, x + i * ix->d, k, ef, tie_rule, vt, T, D + i * k, I + i * k, st);
            s0 += st.ndis; s1 += st.nhops; s2 += st.ndis_up; s3 += st.nhops_up;
            if (per_query) { per_query[2 * i] = (int32_t)st.ndis; per_query[2 * i + 1] = (int32_t)st.nhops; }
        }
    }
    if (stats) { stats[0] = s0; stats[1] = s1; stats[2] = s2; stats[3] = s3; }
    return 0;
}

// Exact top-k under the ADC estimate (scan every code): the recall reference.  Ordered by (dis, id).
int orc_adc_scan(void *h, const float *x, int64_t n, int k, float *D, int64_t *I, int threads)
{
    auto *ix = (OrcIndex *)h;
    if (threads <= 0) {
#ifdef _OPENMP
        threads = omp_get_max_threads();
#else
        threads = 1;
#endif
    }
#pragma omp parallel num_threads(threads)
    {
        std::vector<float> T(ix->M * ix->ksub);
        std::vector<std::pair<float, int64_t>> all((size_t)ix->ntotal);
#pragma omp for schedule(dynamic, 1)
        for (int64_t i = 0; i < n; ++i) {
            adc_table(*ix, x + i * ix->d, T.data());
            for (int64_t v = 0; v < ix->ntotal; ++v) all[(size_t)v] = {adc_dis(*ix, T.data(), v), v};
            int64_t kk = std::min<int64_t>(k, ix->ntotal);
            std::partial_sort(all.begin(), all.begin() + kk, all.end());
            for (int64_t j = 0; j < k; ++j) {
                D[i * k + j] = j < kk ? all[(size_t)j].first : FLT_MAX;
                I[i * k + j] = j < kk ? all[(size_t)j].second : -1;
            }
        }
    }
    return 0;
}

// dis[i] = ADC distance of query x to vertex ids[i] (ids < 0 -> FLT_MAX): lets tests recompute D == dis(I).
void orc_adc_distances(void *h, const float *x, const int64_t *ids, int64_t m, float *dis)
{
    auto *ix = (OrcIndex *)h;
    std::vector<float> T(ix->M * ix->ksub);
    adc_table(*ix, x, T.data());
    for (int64_t i = 0; i < m; ++i) dis[i] = ids[i] < 0 ? FLT_MAX : adc_dis(*ix, T.data(), ids[i]);
}

// Structural checks a FAISS-written graph satisfies; returns 0 or a negative code (tests use it on files
// written by this repo's builder).
int orc_index_validate(void *h)
{
    auto *ix = (OrcIndex *)h;
    if (ix->cum.size() < 2 || ix->cum[0] != 0) return -1;
    for (int64_t v = 0; v < ix->ntotal; ++v) {
        int lv = ix->levels[(size_t)v];
        if (lv < 1 || (size_t)lv >= ix->cum.size()) return -2;
        if (ix->offsets[(size_t)v + 1] - ix->offsets[(size_t)v] != (uint64_t)ix->cum[(size_t)lv]) return -3;
        for (int l = 0; l < lv; ++l) {
            bool ended = false;
            for (uint64_t j = ix->offsets[(size_t)v] + (uint64_t)ix->cum[(size_t)l]; j < ix->offsets[(size_t)v] + (uint64_t)ix->cum[(size_t)l + 1]; ++j) {
                int32_t u = ix->neighbors[j];
                if (u < 0) { ended = true; continue; }
                if (ended) return -4;                        // -1 padding must be a suffix
                if (u >= ix->ntotal || u == v) return -5;
                if (ix->levels[(size_t)u] <= l) return -6;   // neighbour must exist on this level
                for (uint64_t j2 = ix->offsets[(size_t)v] + (uint64_t)ix->cum[(size_t)l]; j2 < j; ++j2)
                    if (ix->neighbors[j2] == u) return -10;  // no vertex twice in a row
            }
        }
    }
    if (ix->ntotal > 0) {
        if (ix->entry_point < 0 || ix->entry_point >= ix->ntotal) return -7;
        if (ix->levels[(size_t)ix->entry_point] != ix->max_level + 1) return -8;
    }
    if (ix->offsets.back() != ix->neighbors.size()) return -9;
    return 0;
}

// Copies of the PQ arrays and graph statistics (builder tests): centroids [M][ksub][dsub], codes ntotal x code_size
// (either may be NULL); deg[0] = mean level-0 out-degree, deg[1] = share of vertices with an empty level-0 row,
// deg[2] = mean out-degree on level 1 over the vertices that have one.
void orc_index_arrays(void *h, float *centroids, uint8_t *codes, double *deg)
{
    auto *ix = (OrcIndex *)h;
    if (centroids) memcpy(centroids, ix->centroids.data(), ix->centroids.size() * sizeof(float));
    if (codes) memcpy(codes, ix->codes.data(), ix->codes.size());
    if (deg) {
        double s0 = 0, empty = 0, s1 = 0, n1 = 0;
        for (int64_t v = 0; v < ix->ntotal; ++v) {
            int c = 0;
            for (int j = 0; j < ix->cum[1]; ++j) c += ix->neighbors[ix->offsets[(size_t)v] + (uint64_t)j] >= 0;
            s0 += c; empty += c == 0;
            if (ix->levels[(size_t)v] > 1) {
                int c1 = 0;
                for (int j = ix->cum[1]; j < ix->cum[2]; ++j) c1 += ix->neighbors[ix->offsets[(size_t)v] + (uint64_t)j] >= 0;
                s1 += c1; n1 += 1;
            }
        }
        deg[0] = s0 / (double)ix->ntotal; deg[1] = empty / (double)ix->ntotal; deg[2] = n1 ? s1 / n1 : 0.0;
    }
}

}  // extern "C"
