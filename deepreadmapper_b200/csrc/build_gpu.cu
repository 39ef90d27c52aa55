// GPU index builder (sm_100a) behind hnswpq_index: product-quantizer training, encoding and HNSW construction on the
// device, written out as a FAISS "IHNp" file.  SURVEY.md §8(f) rank 1; mirrors what the reference asks FAISS to do in
// build_faiss_index (src/hnswpq/index.cpp:86-193): IndexHNSWPQ(dim, M_pq, M_hnsw, nbits); efConstruction = EFC;
// train(sample); add(all).  It is the device counterpart of builder.cpp and shares its seed-derived choices
// (training sample, k-means start, vertex levels, insertion order).
//
//   * PQ training: Lloyd iterations per sub-space on <= 65 536 sampled rows; assignment on the device (centroids of a
//     sub-space in shared memory), per-cluster sums in double with a fixed reduction order (deterministic), centroid
//     update + FAISS's empty-cluster split on the host (256 x dsub numbers per sub-space).
//   * Encoding: nearest centroid per sub-space, first minimum on ties (IndexPQ::add).
//   * HNSW: BATCHED insertion.  Vertices are taken in FAISS's order (descending level, shuffled inside a level) in
//     batches of at most 1/16 of the graph built so far (<= 32 768).  `build_insert_kernel` gives every vertex of the
//     batch one warp: ADC table of its vector, greedy descent to its top level, then on each of its levels the
//     efConstruction-bounded search (the sorted-pool formulation of search_kernel.cu), FAISS's shrink_neighbor_list
//     heuristic on symmetric PQ distances, its own adjacency row, and one "link me back" request per chosen
//     neighbour.  The graph is frozen while a batch searches (a vertex's own rows are unreachable until its back links
//     exist), so the kernel has no races; `build_backlink_kernel` then applies the requests, one warp per request under
//     a per-vertex spin lock (HNSW::add_link: append, or re-shrink the full row).  Batch mates do not see each other,
//     which is the price of parallel insertion (FAISS's OpenMP build has the same kind of window, one vertex per thread).
//   * Adjacency rows are kept FUSED with the PQ codes of the neighbours - the layout of the search image
//     (index_image.cu) - so a construction hop is one contiguous read as well.
#include "common.h"

#include <algorithm>
#include <chrono>
#include <cfloat>
#include <cstdlib>
#include <cstring>

namespace drm {

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int BUILD_WARPS = 4;
typedef unsigned long long u64;

// ------------------------------------------------------------------------------------------------ PQ training
// assign[m][i] = arg min_c || tr[i][m*dsub ..] - cent[m][c] ||^2 (first minimum), plain fp32 mul/add in t order like
// builder.cpp's k-means.  grid = (ceil(nt / 256), M); the sub-space's 256 centroids sit in shared memory.
__global__ void __launch_bounds__(256) pq_assign_kernel(const float *__restrict__ rows, long long n, int d, int dsub,
                                                         const float *__restrict__ cent, int *__restrict__ assign)
{
    extern __shared__ float sc[];  // 256 x dsub
    const int m = blockIdx.y;
    for (int i = threadIdx.x; i < 256 * dsub; i += blockDim.x) sc[i] = cent[(size_t)m * 256 * dsub + i];
    __syncthreads();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *xi = rows + (size_t)i * d + (size_t)m * dsub;
    float best = FLT_MAX;
    int bi = 0;
    for (int c = 0; c < 256; ++c) {
        float s = 0.f;
        for (int t = 0; t < dsub; ++t) {
            const float e = __fsub_rn(__ldg(xi + t), sc[c * dsub + t]);
            s = __fadd_rn(s, __fmul_rn(e, e));
        }
        if (s < best) { best = s; bi = c; }
    }
    assign[(size_t)m * n + i] = bi;
}

// Encoding writes bytes: codes[i][m].
__global__ void __launch_bounds__(256) pq_encode_kernel(const float *__restrict__ x, long long n, int d, int dsub, int M,
                                                         const float *__restrict__ cent, uint8_t *__restrict__ codes)
{
    extern __shared__ float sc[];
    const int m = blockIdx.y;
    for (int i = threadIdx.x; i < 256 * dsub; i += blockDim.x) sc[i] = cent[(size_t)m * 256 * dsub + i];
    __syncthreads();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float *xi = x + (size_t)i * d + (size_t)m * dsub;
        float best = FLT_MAX;
        int bj = 0;
        for (int c = 0; c < 256; ++c) {
            float s = 0.f;
            for (int t = 0; t < dsub; ++t) {
                const float e = __fsub_rn(__ldg(xi + t), sc[c * dsub + t]);
                s = __fadd_rn(s, __fmul_rn(e, e));
            }
            if (s < best) { best = s; bj = c; }
        }
        codes[(size_t)i * M + m] = (uint8_t)bj;
    }
}

// Per-cluster sums: one CTA per (cluster c, sub-space m); thread t accumulates rows t, t+256, ... in double, then a
// fixed-shape tree adds the 256 partials - the result does not depend on scheduling.
__global__ void __launch_bounds__(256) pq_sum_kernel(const float *__restrict__ rows, long long n, int d, int dsub,
                                                      const int *__restrict__ assign, double *__restrict__ sum,
                                                      long long *__restrict__ cnt)
{
    extern __shared__ double sd[];  // 256 x (dsub + 1)
    const int c = blockIdx.x, m = blockIdx.y, tid = threadIdx.x;
    double *mine = sd + (size_t)tid * (dsub + 1);
    for (int t = 0; t <= dsub; ++t) mine[t] = 0.0;
    const int *as = assign + (size_t)m * n;
    for (long long i = tid; i < n; i += 256)
        if (as[i] == c) {
            const float *xi = rows + (size_t)i * d + (size_t)m * dsub;
            for (int t = 0; t < dsub; ++t) mine[t] += (double)xi[t];
            mine[dsub] += 1.0;
        }
    __syncthreads();
    for (int step = 128; step > 0; step >>= 1) {
        if (tid < step) {
            double *o = sd + (size_t)(tid + step) * (dsub + 1);
            for (int t = 0; t <= dsub; ++t) mine[t] += o[t];
        }
        __syncthreads();
    }
    if (tid == 0) {
        for (int t = 0; t < dsub; ++t) sum[((size_t)m * 256 + c) * dsub + t] = mine[t];
        cnt[(size_t)m * 256 + c] = (long long)mine[dsub];
    }
}

// gather of the training rows: out[i] = x[pick[i] * step]
__global__ void gather_rows_kernel(const float *__restrict__ x, const long long *__restrict__ pick, long long step, long long nt,
                                   int d, float *__restrict__ out)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nt * d) return;
    const long long i = t / d;
    const int c = (int)(t % d);
    out[t] = x[(size_t)(pick[i] * step) * d + c];
}

// sdc[m][a][b] = || cent[m][a] - cent[m][b] ||^2 (ProductQuantizer::compute_sdc_table)
__global__ void sdc_kernel(const float *__restrict__ cent, int M, int dsub, float *__restrict__ sdc)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= M * 256 * 256) return;
    const int m = t >> 16, a = (t >> 8) & 255, b = t & 255;
    const float *ca = cent + ((size_t)m * 256 + a) * dsub, *cb = cent + ((size_t)m * 256 + b) * dsub;
    float s = 0.f;
    for (int k = 0; k < dsub; ++k) {
        const float e = __fsub_rn(ca[k], cb[k]);
        s = __fadd_rn(s, __fmul_rn(e, e));
    }
    sdc[t] = s;
}

// ------------------------------------------------------------------------------------------------ HNSW construction
struct GraphParams {
    const float *x;           // n x d
    const int *order;         // insertion order
    const float *centroids;   // [M][256][dsub]
    const uint8_t *codes;     // n x CS
    const int *levels;        // top level + 1
    const int *up_base;       // first upper row of a vertex (its level-1 row), -1 when it has none
    uint8_t *rows0, *up;      // fused rows: [W ids][W codes]
    int row0_bytes, rowu_bytes, W0, Wu, c0_off, cu_off;
    const float *sdc;         // [M][256][256]
    int *locks;
    int4 *requests;           // (target, source, level, -)
    unsigned *ctr;            // this batch: [0] insert work counter, [1] number of requests, [2] back-link work counter
    int d, M, dsub, efc, cap_alloc;
    int entry, max_level;
    long long b0, b1;         // positions of the batch in `order`
    u64 *stats;               // [0] hops, [1] ADC evaluations, [2] appended links, [3] re-shrunk rows, [4] lock spins
};

template <int CS> struct CodeW { unsigned w[CS / 4]; };

template <int CS> __device__ __forceinline__ CodeW<CS> ld_code_nc(const uint8_t *p);
template <> __device__ __forceinline__ CodeW<8> ld_code_nc<8>(const uint8_t *p)
{
    const uint2 v = __ldg(reinterpret_cast<const uint2 *>(p));
    CodeW<8> c; c.w[0] = v.x; c.w[1] = v.y; return c;
}
template <> __device__ __forceinline__ CodeW<16> ld_code_nc<16>(const uint8_t *p)
{
    const uint4 v = __ldg(reinterpret_cast<const uint4 *>(p));
    CodeW<16> c; c.w[0] = v.x; c.w[1] = v.y; c.w[2] = v.z; c.w[3] = v.w; return c;
}
template <int CS> __device__ __forceinline__ CodeW<CS> ld_code_cg(const uint8_t *p);
template <> __device__ __forceinline__ CodeW<8> ld_code_cg<8>(const uint8_t *p)
{
    const uint2 v = __ldcg(reinterpret_cast<const uint2 *>(p));
    CodeW<8> c; c.w[0] = v.x; c.w[1] = v.y; return c;
}
template <> __device__ __forceinline__ CodeW<16> ld_code_cg<16>(const uint8_t *p)
{
    const uint4 v = __ldcg(reinterpret_cast<const uint4 *>(p));
    CodeW<16> c; c.w[0] = v.x; c.w[1] = v.y; c.w[2] = v.z; c.w[3] = v.w; return c;
}
template <int CS> __device__ __forceinline__ void st_code(uint8_t *p, const CodeW<CS> &c);
template <> __device__ __forceinline__ void st_code<8>(uint8_t *p, const CodeW<8> &c)
{
    *reinterpret_cast<uint2 *>(p) = make_uint2(c.w[0], c.w[1]);
}
template <> __device__ __forceinline__ void st_code<16>(uint8_t *p, const CodeW<16> &c)
{
    *reinterpret_cast<uint4 *>(p) = make_uint4(c.w[0], c.w[1], c.w[2], c.w[3]);
}
template <int CS> __device__ __forceinline__ CodeW<CS> shfl_code(const CodeW<CS> &c, int src)
{
    CodeW<CS> o;
#pragma unroll
    for (int i = 0; i < CS / 4; ++i) o.w[i] = __shfl_sync(FULL, c.w[i], src);
    return o;
}
template <int CS> __device__ __forceinline__ CodeW<CS> zero_code()
{
    CodeW<CS> o;
#pragma unroll
    for (int i = 0; i < CS / 4; ++i) o.w[i] = 0u;
    return o;
}
// ADC distance: sum_m T[m][code[m]], m ascending (same order as the search kernel and the oracle).
template <int CS> __device__ __forceinline__ float adc_code(const float *T, const CodeW<CS> &c)
{
    float acc = 0.f;
#pragma unroll
    for (int m = 0; m < CS; ++m) {
        const float t = T[m * 256 + ((c.w[m >> 2] >> (8 * (m & 3))) & 0xffu)];
        acc = m == 0 ? t : __fadd_rn(acc, t);
    }
    return acc;
}
// symmetric PQ distance between two codes (PQDistanceComputer::symmetric_dis): sum_m sdc[m][a[m]][b[m]]
template <int CS> __device__ __forceinline__ float sym_code(const float *__restrict__ sdc, const CodeW<CS> &a, const CodeW<CS> &b)
{
    float t[CS];
#pragma unroll
    for (int m = 0; m < CS; ++m) {
        const unsigned ca = (a.w[m >> 2] >> (8 * (m & 3))) & 0xffu, cb = (b.w[m >> 2] >> (8 * (m & 3))) & 0xffu;
        t[m] = __ldg(sdc + (((size_t)m * 256 + ca) << 8) + cb);
    }
    float acc = t[0];
#pragma unroll
    for (int m = 1; m < CS; ++m) acc = __fadd_rn(acc, t[m]);
    return acc;
}

__device__ __forceinline__ u64 key_of(float d, int id) { return ((u64)__float_as_uint(d) << 32) | ((u64)(unsigned)id << 1); }
__device__ __forceinline__ int key_id(u64 k) { return (int)((unsigned)(k & 0xffffffffu) >> 1); }
__device__ __forceinline__ float key_dis(u64 k) { return __uint_as_float((unsigned)(k >> 32)); }

__device__ __forceinline__ u64 warp_min64(u64 v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const u64 other = __shfl_xor_sync(FULL, v, o);
        v = other < v ? other : v;
    }
    return v;
}

__device__ __forceinline__ uint8_t *row_of(const GraphParams &p, int v, int level, int &W, int &coff)
{
    if (level == 0) { W = p.W0; coff = p.c0_off; return p.rows0 + (size_t)v * p.row0_bytes; }
    W = p.Wu; coff = p.cu_off;
    return p.up + (size_t)(__ldg(p.up_base + v) + level - 1) * p.rowu_bytes;
}

// T[m][j] = sum_t (x[m*dsub+t] - c[m][j][t])^2, one fma per term in t order (the search kernel's table).
__device__ __forceinline__ void build_table(float *T, const float *__restrict__ x, const float *__restrict__ cent, int M, int dsub, int lane)
{
    for (int m = 0; m < M; ++m) {
        const float *xm = x + m * dsub;
        for (int j = lane; j < 256; j += 32) {
            const float *c = cent + ((size_t)(m * 256 + j)) * dsub;
            float acc = 0.f;
            for (int t = 0; t < dsub; ++t) {
                const float e = __fsub_rn(__ldg(xm + t), __ldg(c + t));
                acc = __fmaf_rn(e, e, acc);
            }
            T[m * 256 + j] = acc;
        }
    }
}

// One warp per new vertex: search + neighbour selection + own rows + back-link requests.  The graph is read-only here.
template <int CS>
__global__ void __launch_bounds__(BUILD_WARPS * 32) build_insert_kernel(const GraphParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t per_warp = (size_t)p.M * 256 * sizeof(float) + (size_t)p.cap_alloc * sizeof(u64);
    float *T = reinterpret_cast<float *>(smem_raw + warp * per_warp);
    u64 *pool = reinterpret_cast<u64 *>(smem_raw + warp * per_warp + (size_t)p.M * 256 * sizeof(float));
    const int cap = p.efc;
    unsigned nhops = 0, ndis = 0;

    for (;;) {
        unsigned wi = 0;
        if (lane == 0) wi = atomicAdd(p.ctr + 0, 1u);
        wi = __shfl_sync(FULL, wi, 0);
        const long long idx = p.b0 + (long long)wi;
        if (idx >= p.b1) break;
        const int pt = __ldg(p.order + idx);
        const int pt_level = __ldg(p.levels + pt) - 1;

        __syncwarp();
        build_table(T, p.x + (size_t)pt * p.d, p.centroids, p.M, p.dsub, lane);
        __syncwarp();

        int nearest = p.entry;
        float d_near = adc_code<CS>(T, ld_code_nc<CS>(p.codes + (size_t)nearest * CS));
        int level = p.max_level;

        // ---- greedy descent above the vertex's top level (HNSW::add_with_locks -> greedy_update_nearest)
        for (; level > pt_level; --level) {
            for (;;) {
                int W, coff;
                const uint8_t *row = row_of(p, nearest, level, W, coff);
                const int nid = lane < W ? __ldg(reinterpret_cast<const int *>(row) + lane) : -1;
                const unsigned neg = __ballot_sync(FULL, nid < 0);
                const int first_neg = neg ? (__ffs(neg) - 1) : 32;
                float dj = __uint_as_float(0x7f800000u);
                if (lane < first_neg) dj = adc_code<CS>(T, ld_code_nc<CS>(row + coff + CS * lane));
                const u64 best = warp_min64(((u64)__float_as_uint(dj) << 32) | (unsigned)lane);
                const float dmin = __uint_as_float((unsigned)(best >> 32));
                ++nhops;
                if (dmin < d_near) {
                    d_near = dmin;
                    nearest = __shfl_sync(FULL, nid, (int)(best & 31u));
                } else break;
            }
        }

        // ---- the vertex's own levels, top down (add_links_starting_from)
        for (; level >= 0; --level) {
            // efConstruction-bounded search on this level: sorted pool of (distance, id) keys, bit 0 = expanded
            int size = 1, cur = 0;
            if (lane == 0) pool[0] = key_of(d_near, nearest);
            __syncwarp();
            while (cur >= 0) {
                const u64 k0 = pool[cur];
                int W, coff;
                const uint8_t *row = row_of(p, key_id(k0), level, W, coff);
                const int nid = lane < W ? __ldg(reinterpret_cast<const int *>(row) + lane) : -1;
                CodeW<CS> code = zero_code<CS>();
                if (lane < W) code = ld_code_nc<CS>(row + coff + CS * lane);
                __syncwarp();
                if (lane == 0) pool[cur] = k0 | 1ull;
                __syncwarp();
                ++nhops;
                const unsigned neg = __ballot_sync(FULL, nid < 0);
                const int first_neg = neg ? (__ffs(neg) - 1) : 32;
                bool cand = lane < first_neg;
                ndis += first_neg;
                u64 key = 0;
                if (cand) key = key_of(adc_code<CS>(T, code), nid);
                const u64 maxkey = size == cap ? (pool[cap - 1] & ~1ull) : ~0ull;
                cand = cand && key < maxkey;
                int lo = 0;
                {
                    int l = 0, h = size;
                    const int iters = 32 - __clz(size);
                    for (int it = 0; it < iters; ++it)
                        if (l < h) {
                            const int mid = (l + h) >> 1;
                            if ((pool[mid] & ~1ull) < key) l = mid + 1; else h = mid;
                        }
                    lo = l;
                    if (cand && l < size && (pool[l] & ~1ull) == key) cand = false;  // already in the pool
                }
                int hint = cur + 1;
                unsigned mask = __ballot_sync(FULL, cand);
                while (mask) {
                    const int src = __ffs(mask) - 1;
                    mask &= mask - 1;
                    const u64 kin = __shfl_sync(FULL, key, src);
                    const int pin = __shfl_sync(FULL, lo, src);
                    if (pin >= cap) continue;
                    if (pin < size && (pool[pin] & ~1ull) == kin) continue;  // the same id twice in one row
                    const int nsz = size < cap ? size + 1 : cap;
                    for (int top = nsz; top > pin + 1; top -= 32) {
                        const int i = top - 1 - lane;
                        u64 tmp = 0;
                        if (i > pin) tmp = pool[i - 1];
                        __syncwarp();
                        if (i > pin) pool[i] = tmp;
                        __syncwarp();
                    }
                    if (lane == 0) pool[pin] = kin;
                    __syncwarp();
                    size = nsz;
                    if (key > kin) lo += 1;
                    hint = pin < hint ? pin : hint;
                }
                cur = -1;
                for (int b = hint; b < size; b += 32) {
                    const int i = b + lane;
                    const unsigned m = __ballot_sync(FULL, i < size && !(pool[i] & 1ull));
                    if (m) { cur = b + __ffs(m) - 1; break; }
                }
            }

            // shrink_neighbor_list: walk the pool in ascending distance, keep v1 unless a kept vertex is closer to it than
            // the new vertex is (symmetric PQ distance vs ADC distance); lane j holds kept vertex j.
            const int width = level == 0 ? p.W0 : p.Wu;
            int nch = 0, myid = -1;
            CodeW<CS> mycode = zero_code<CS>();
            for (int i0 = 0; i0 < size && nch < width; i0 += 32) {
                const int i = i0 + lane;
                const u64 kk = i < size ? pool[i] : 0ull;
                CodeW<CS> cc = zero_code<CS>();
                if (i < size) cc = ld_code_nc<CS>(p.codes + (size_t)key_id(kk) * CS);
                const int lim = size - i0 < 32 ? size - i0 : 32;
                for (int j = 0; j < lim && nch < width; ++j) {
                    const u64 k1 = __shfl_sync(FULL, kk, j);
                    const CodeW<CS> c1 = shfl_code<CS>(cc, j);
                    const float d1 = key_dis(k1);
                    const bool bad = lane < nch && sym_code<CS>(p.sdc, c1, mycode) < d1;
                    if (!__any_sync(FULL, bad)) {
                        if (lane == nch) { myid = key_id(k1); mycode = c1; }
                        ++nch;
                    }
                }
            }
            // own row (nobody can reach it yet) and the back-link requests
            {
                int W, coff;
                uint8_t *row = row_of(p, pt, level, W, coff);
                if (lane < W) {
                    reinterpret_cast<int *>(row)[lane] = lane < nch ? myid : -1;
                    st_code<CS>(row + coff + CS * lane, lane < nch ? mycode : zero_code<CS>());
                }
                unsigned base = 0;
                if (lane == 0) base = atomicAdd(p.ctr + 1, (unsigned)nch);
                base = __shfl_sync(FULL, base, 0);
                if (lane < nch) p.requests[base + lane] = make_int4(myid, pt, level, 0);
            }
            // next level starts from the closest vertex found on this one
            const u64 kbest = pool[0];
            nearest = key_id(kbest);
            d_near = key_dis(kbest);
            __syncwarp();
        }
    }
    if (p.stats && lane == 0) {
        atomicAdd(p.stats + 0, (u64)nhops);
        atomicAdd(p.stats + 1, (u64)ndis);
    }
}

// One warp per back-link request (target <- source on `level`), HNSW::add_link under the target's lock: a free slot
// takes the source; a full row is re-selected from its W entries + the source by shrink_neighbor_list on symmetric
// distances to the target.  Rows are read with ld.global.cg and published with a fence before the lock is released.
template <int CS>
__global__ void __launch_bounds__(BUILD_WARPS * 32) build_backlink_kernel(const GraphParams p)
{
    const int lane = threadIdx.x & 31;
    const unsigned nreq = __ldcg(p.ctr + 1);
    unsigned n_append = 0, n_shrink = 0, n_spin = 0;
    for (;;) {
        unsigned ri = 0;
        if (lane == 0) ri = atomicAdd(p.ctr + 2, 1u);
        ri = __shfl_sync(FULL, ri, 0);
        if (ri >= nreq) break;
        const int4 rq = p.requests[ri];
        const int tgt = rq.x, src = rq.y, level = rq.z;
        int W, coff;
        uint8_t *row = row_of(p, tgt, level, W, coff);
        const CodeW<CS> csrc = ld_code_nc<CS>(p.codes + (size_t)src * CS);
        const CodeW<CS> ctgt = ld_code_nc<CS>(p.codes + (size_t)tgt * CS);

        if (lane == 0) {
            while (atomicCAS(p.locks + tgt, 0, 1) != 0) { ++n_spin; __nanosleep(64); }
            __threadfence();
        }
        __syncwarp();

        const int id = lane < W ? __ldcg(reinterpret_cast<const int *>(row) + lane) : -2;
        if (!__any_sync(FULL, id == src)) {
            const unsigned freem = __ballot_sync(FULL, id == -1);
            if (freem) {
                const int slot = __ffs(freem) - 1;
                if (lane == slot) {
                    st_code<CS>(row + coff + CS * slot, csrc);
                    reinterpret_cast<int *>(row)[slot] = src;
                }
                ++n_append;
            } else {
                ++n_shrink;
                CodeW<CS> cj = zero_code<CS>();
                float dj = __uint_as_float(0x7f800000u);
                u64 kj = ~0ull;
                if (lane < W) {
                    cj = ld_code_cg<CS>(row + coff + CS * lane);
                    dj = sym_code<CS>(p.sdc, ctgt, cj);
                    kj = ((u64)__float_as_uint(dj) << 32) | (unsigned)id;
                }
                const float dsrc = sym_code<CS>(p.sdc, ctgt, csrc);
                const u64 ksrc = ((u64)__float_as_uint(dsrc) << 32) | (unsigned)src;
                // rank of every candidate among the W + 1 (keys are distinct: ids are)
                int rank = ksrc < kj ? 1 : 0;
                for (int j = 0; j < W; ++j) {
                    const u64 o = __shfl_sync(FULL, kj, j);
                    rank += o < kj ? 1 : 0;
                }
                const int rank_src = __popc(__ballot_sync(FULL, lane < W && kj < ksrc));
                bool chosen = false, src_chosen = false;
                int pos = -1, pos_src = -1, nch = 0;
                for (int r = 0; r <= W && nch < W; ++r) {
                    const unsigned own = __ballot_sync(FULL, lane < W && rank == r);
                    const int o = own ? __ffs(own) - 1 : 0;
                    const bool is_src = r == rank_src;
                    CodeW<CS> c1 = shfl_code<CS>(cj, o);
                    float d1 = __shfl_sync(FULL, dj, o);
                    if (is_src) { c1 = csrc; d1 = dsrc; }
                    bool bad = chosen && sym_code<CS>(p.sdc, c1, cj) < d1;
                    if (src_chosen && lane == 0 && sym_code<CS>(p.sdc, c1, csrc) < d1) bad = true;
                    if (!__any_sync(FULL, bad)) {
                        if (is_src) { src_chosen = true; pos_src = nch; }
                        else if (lane == o) { chosen = true; pos = nch; }
                        ++nch;
                    }
                }
                __syncwarp();
                // every slot has exactly one writer: kept entries go to their selection position, the rest become -1
                if (lane < W && lane >= nch) {
                    reinterpret_cast<int *>(row)[lane] = -1;
                    st_code<CS>(row + coff + CS * lane, zero_code<CS>());
                }
                if (chosen) {
                    reinterpret_cast<int *>(row)[pos] = id;
                    st_code<CS>(row + coff + CS * pos, cj);
                }
                if (src_chosen && lane == 0) {
                    reinterpret_cast<int *>(row)[pos_src] = src;
                    st_code<CS>(row + coff + CS * pos_src, csrc);
                }
            }
        }
        __threadfence();
        __syncwarp();
        if (lane == 0) atomicExch(p.locks + tgt, 0);
    }
    if (p.stats && lane == 0) {
        atomicAdd(p.stats + 2, (u64)n_append);
        atomicAdd(p.stats + 3, (u64)n_shrink);
        atomicAdd(p.stats + 4, (u64)n_spin);
    }
}

// rows start empty: ids -1, codes 0
__global__ void init_rows_kernel(uint8_t *rows, long long n_rows, int row_bytes, int W)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int words = row_bytes / 4;
    if (t >= n_rows * words) return;
    const int w = (int)(t % words);
    reinterpret_cast<int *>(rows)[t] = w < W ? -1 : 0;
}

// fused rows -> FAISS `neighbors` (ids only): level 0, then the upper rows
__global__ void export_rows0_kernel(const uint8_t *__restrict__ rows0, long long n, int row_bytes, int W0,
                                    const unsigned long long *__restrict__ offsets, int *__restrict__ neighbors)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * W0) return;
    const long long v = t / W0;
    const int slot = (int)(t % W0);
    neighbors[offsets[v] + slot] = reinterpret_cast<const int *>(rows0 + (size_t)v * row_bytes)[slot];
}
__global__ void export_up_kernel(const uint8_t *__restrict__ up, long long n_rows, int row_bytes, int Wu, int W0,
                                 const int *__restrict__ row_vertex, const int *__restrict__ row_level,
                                 const unsigned long long *__restrict__ offsets, int *__restrict__ neighbors)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_rows * Wu) return;
    const long long r = t / Wu;
    const int slot = (int)(t % Wu);
    const int v = row_vertex[r], l = row_level[r];
    neighbors[offsets[v] + W0 + (size_t)(l - 1) * Wu + slot] = reinterpret_cast<const int *>(up + (size_t)r * row_bytes)[slot];
}

inline int round16i(int x) { return (x + 15) / 16 * 16; }

struct Dev {  // device allocations of one build, freed together
    std::vector<void *> ptrs;
    ~Dev() { for (void *q : ptrs) cudaFree(q); }
    template <class T> cudaError_t alloc(T **p, size_t bytes)
    {
        void *q = nullptr;
        cudaError_t e = cudaMalloc(&q, bytes ? bytes : 16);
        if (e == cudaSuccess) ptrs.push_back(q);
        *p = static_cast<T *>(q);
        return e;
    }
};

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

// x: n x d (host, or device when x_on_device).  Training rows: `train` (host, n_train x d) when given, else rows
// 0, train_step, 2*train_step, ... of x (the reference trains on an evenly strided 50 % sample, src/hnswpq/index.cpp:57-84,99).
// timing_s (nullable, 5 doubles): upload, PQ training, encoding, graph construction, export.
int build_index_gpu(const float *x, bool x_on_device, int64_t n, int d, const float *train, int64_t n_train, int64_t train_step,
                    int M_pq, int nbits, int M_hnsw, int efc, uint32_t seed, int device, HostIndex &ix, double *timing_s)
{
    if (!x || n <= 0 || d <= 0 || M_pq <= 0 || d % M_pq) return fail(DRM_EINVAL, "bad index build arguments (d must be a multiple of M_pq)");
    if (nbits != 8) return fail(DRM_EUNSUPPORTED, "only nbits == 8 is supported by the builder");
    if (M_hnsw < 2 || efc < 1) return fail(DRM_EINVAL, "M_hnsw must be >= 2 and EFC >= 1");
    if (M_hnsw > 16) return fail(DRM_EUNSUPPORTED, "the GPU builder handles M_hnsw <= 16 (one level-0 row per warp pass)");
    if (M_pq != 8 && M_pq != 16) return fail(DRM_EUNSUPPORTED, "the GPU builder handles M_pq == 8 or 16");
    if (efc > 2048) return fail(DRM_EUNSUPPORTED, "the GPU builder handles EFC <= 2048");
    if (n >= ((int64_t)1 << 30)) return fail(DRM_EUNSUPPORTED, "more than 2^30 vectors in one build");
    if (train_step < 1) train_step = 1;
    if (train && x_on_device) return fail(DRM_EINVAL, "a device-resident build takes its training rows from x (train must be NULL)");
    const int dsub = d / M_pq, CS = M_pq;
    if (dsub > 96) return fail(DRM_EUNSUPPORTED, "sub-vector dimension above 96");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return fail(DRM_ECUDA, "no CUDA device: the GPU index builder has no CPU fallback"); }
    DRM_CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    DRM_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    const bool verbose = getenv("DRM_BUILD_VERBOSE") != nullptr;
    double t_mark = now_s();
    double tim[5] = {0, 0, 0, 0, 0};
    auto lap = [&](int slot) { cudaDeviceSynchronize(); const double t = now_s(); tim[slot] += t - t_mark; t_mark = t; };

    Dev mem;
    ix = HostIndex();
    ix.d = d; ix.ntotal = n; ix.M = (uint64_t)M_pq; ix.nbits = 8; ix.ksub = 256; ix.dsub = (uint64_t)dsub; ix.code_size = (uint64_t)CS;
    ix.efConstruction = efc; ix.efSearch = 16; ix.upper_beam = 1;

    // ---- vectors on the device
    const float *dx = x;
    if (!x_on_device) {
        float *buf = nullptr;
        DRM_CUDA_TRY(mem.alloc(&buf, (size_t)n * d * sizeof(float)));
        DRM_CUDA_TRY(cudaMemcpy(buf, x, (size_t)n * d * sizeof(float), cudaMemcpyHostToDevice));
        dx = buf;
    }
    lap(0);

    // ---- ProductQuantizer::train
    if (!train) n_train = (n + train_step - 1) / train_step;
    else train_step = 1;
    std::vector<int64_t> pick;
    builder_training_pick(n_train, seed, pick);
    const int64_t nt = (int64_t)pick.size();
    std::vector<float> tr((size_t)nt * d);
    float *d_tr = nullptr;
    DRM_CUDA_TRY(mem.alloc(&d_tr, tr.size() * sizeof(float)));
    if (x_on_device) {
        long long *d_pick = nullptr;
        DRM_CUDA_TRY(mem.alloc(&d_pick, (size_t)nt * sizeof(long long)));
        std::vector<long long> pk(pick.begin(), pick.end());
        DRM_CUDA_TRY(cudaMemcpy(d_pick, pk.data(), (size_t)nt * sizeof(long long), cudaMemcpyHostToDevice));
        const long long th = (long long)nt * d;
        gather_rows_kernel<<<(unsigned)((th + 255) / 256), 256>>>(dx, d_pick, train_step, nt, d, d_tr);
        DRM_CUDA_TRY(cudaGetLastError());
        DRM_CUDA_TRY(cudaMemcpy(tr.data(), d_tr, tr.size() * sizeof(float), cudaMemcpyDeviceToHost));
    } else {
        const float *src = train ? train : x;
        for (int64_t i = 0; i < nt; ++i) memcpy(&tr[(size_t)i * d], src + (size_t)(pick[(size_t)i] * train_step) * d, sizeof(float) * d);
        DRM_CUDA_TRY(cudaMemcpy(d_tr, tr.data(), tr.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    ix.centroids.assign((size_t)M_pq * 256 * dsub, 0.f);
    for (int m = 0; m < M_pq; ++m) {
        std::vector<int64_t> start;
        builder_kmeans_init_rows(nt, 256, seed + 1 + (uint32_t)m, start);
        for (int c = 0; c < 256; ++c)
            memcpy(&ix.centroids[((size_t)m * 256 + c) * dsub], &tr[(size_t)start[(size_t)c] * d + (size_t)m * dsub], sizeof(float) * dsub);
    }
    float *d_cent = nullptr;
    int *d_assign = nullptr;
    double *d_sum = nullptr;
    long long *d_cnt = nullptr;
    DRM_CUDA_TRY(mem.alloc(&d_cent, ix.centroids.size() * sizeof(float)));
    DRM_CUDA_TRY(mem.alloc(&d_assign, (size_t)M_pq * nt * sizeof(int)));
    DRM_CUDA_TRY(mem.alloc(&d_sum, (size_t)M_pq * 256 * dsub * sizeof(double)));
    DRM_CUDA_TRY(mem.alloc(&d_cnt, (size_t)M_pq * 256 * sizeof(long long)));
    const size_t smem_assign = (size_t)256 * dsub * sizeof(float), smem_sum = (size_t)256 * (dsub + 1) * sizeof(double);
    DRM_CUDA_TRY(cudaFuncSetAttribute(pq_assign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_assign));
    DRM_CUDA_TRY(cudaFuncSetAttribute(pq_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_assign));
    DRM_CUDA_TRY(cudaFuncSetAttribute(pq_sum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_sum));
    {
        std::vector<double> sum((size_t)M_pq * 256 * dsub);
        std::vector<long long> cnt_ll((size_t)M_pq * 256);
        std::vector<int64_t> cnt((size_t)256);
        for (int it = 0; it < 25; ++it) {
            DRM_CUDA_TRY(cudaMemcpy(d_cent, ix.centroids.data(), ix.centroids.size() * sizeof(float), cudaMemcpyHostToDevice));
            pq_assign_kernel<<<dim3((unsigned)((nt + 255) / 256), (unsigned)M_pq), 256, smem_assign>>>(d_tr, nt, d, dsub, d_cent, d_assign);
            pq_sum_kernel<<<dim3(256, (unsigned)M_pq), 256, smem_sum>>>(d_tr, nt, d, dsub, d_assign, d_sum, d_cnt);
            DRM_CUDA_TRY(cudaGetLastError());
            DRM_CUDA_TRY(cudaMemcpy(sum.data(), d_sum, sum.size() * sizeof(double), cudaMemcpyDeviceToHost));
            DRM_CUDA_TRY(cudaMemcpy(cnt_ll.data(), d_cnt, cnt_ll.size() * sizeof(long long), cudaMemcpyDeviceToHost));
            for (int m = 0; m < M_pq; ++m) {
                for (int c = 0; c < 256; ++c) cnt[(size_t)c] = (int64_t)cnt_ll[(size_t)m * 256 + c];
                builder_kmeans_update(&sum[(size_t)m * 256 * dsub], cnt.data(), 256, dsub, &ix.centroids[(size_t)m * 256 * dsub]);
            }
        }
        DRM_CUDA_TRY(cudaMemcpy(d_cent, ix.centroids.data(), ix.centroids.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    lap(1);

    // ---- IndexPQ::add
    uint8_t *d_codes = nullptr;
    DRM_CUDA_TRY(mem.alloc(&d_codes, (size_t)n * CS));
    {
        const long long blocks = std::min<long long>((n + 255) / 256, (long long)prop.multiProcessorCount * 16);
        pq_encode_kernel<<<dim3((unsigned)blocks, (unsigned)M_pq), 256, smem_assign>>>(dx, n, d, dsub, M_pq, d_cent, d_codes);
        DRM_CUDA_TRY(cudaGetLastError());
    }
    ix.codes.resize((size_t)n * CS);
    DRM_CUDA_TRY(cudaMemcpy(ix.codes.data(), d_codes, ix.codes.size(), cudaMemcpyDeviceToHost));
    lap(2);

    // ---- HNSW: levels, insertion order, row numbering of the upper levels (host), rows on the device
    builder_level_law(ix, n, M_hnsw, seed);
    std::vector<int64_t> order64;
    builder_insertion_order(ix, seed, order64);
    std::vector<int32_t> order(order64.begin(), order64.end());
    const int W0 = 2 * M_hnsw, Wu = M_hnsw;
    GraphParams p;
    memset(&p, 0, sizeof p);
    p.W0 = W0; p.Wu = Wu;
    p.c0_off = round16i(4 * W0); p.row0_bytes = round16i(p.c0_off + CS * W0);
    p.cu_off = round16i(4 * Wu); p.rowu_bytes = round16i(p.cu_off + CS * Wu);
    std::vector<int32_t> up_base((size_t)n, -1), row_vertex, row_level;
    int64_t n_up = 0;
    for (int64_t v = 0; v < n; ++v) {
        const int lv = ix.levels[(size_t)v];
        if (lv > 1) {
            up_base[(size_t)v] = (int32_t)n_up;
            for (int l = 1; l < lv; ++l) { row_vertex.push_back((int32_t)v); row_level.push_back(l); }
            n_up += lv - 1;
        }
    }
    const int entry = order[0], max_level = ix.levels[(size_t)entry] - 1;

    // batches: at most 1/ratio of the graph built so far, capped
    int64_t batch_cap = 32768, ratio = 16;
    if (const char *e = getenv("DRM_BUILD_BATCH")) batch_cap = std::max<int64_t>(1, atoll(e));
    if (const char *e = getenv("DRM_BUILD_RATIO")) ratio = std::max<int64_t>(1, atoll(e));
    std::vector<std::pair<int64_t, int64_t>> batches;
    size_t max_req = 1;
    for (int64_t b0 = 1; b0 < n;) {
        const int64_t b1 = std::min<int64_t>(n, b0 + std::max<int64_t>(1, std::min<int64_t>(batch_cap, b0 / ratio)));
        size_t req = 0;
        for (int64_t i = b0; i < b1; ++i) req += (size_t)ix.cum[(size_t)ix.levels[(size_t)order[(size_t)i]]];
        max_req = std::max(max_req, req);
        batches.emplace_back(b0, b1);
        b0 = b1;
    }

    int *d_levels = nullptr, *d_order = nullptr, *d_upb = nullptr, *d_rv = nullptr, *d_rl = nullptr, *d_locks = nullptr, *d_nb = nullptr;
    unsigned long long *d_off = nullptr;
    float *d_sdc = nullptr;
    int4 *d_req = nullptr;
    unsigned *d_ctr = nullptr;
    u64 *d_stats = nullptr;
    DRM_CUDA_TRY(mem.alloc(&d_levels, (size_t)n * sizeof(int)));
    DRM_CUDA_TRY(mem.alloc(&d_order, (size_t)n * sizeof(int)));
    DRM_CUDA_TRY(mem.alloc(&d_upb, (size_t)n * sizeof(int)));
    DRM_CUDA_TRY(mem.alloc(&d_rv, (size_t)std::max<int64_t>(n_up, 1) * sizeof(int)));
    DRM_CUDA_TRY(mem.alloc(&d_rl, (size_t)std::max<int64_t>(n_up, 1) * sizeof(int)));
    DRM_CUDA_TRY(mem.alloc(&d_locks, (size_t)n * sizeof(int)));
    DRM_CUDA_TRY(mem.alloc(&p.rows0, (size_t)n * p.row0_bytes));
    DRM_CUDA_TRY(mem.alloc(&p.up, (size_t)std::max<int64_t>(n_up, 1) * p.rowu_bytes));
    DRM_CUDA_TRY(mem.alloc(&d_sdc, (size_t)M_pq * 65536 * sizeof(float)));
    DRM_CUDA_TRY(mem.alloc(&d_req, max_req * sizeof(int4)));
    DRM_CUDA_TRY(mem.alloc(&d_ctr, (batches.size() + 1) * 4 * sizeof(unsigned)));
    DRM_CUDA_TRY(mem.alloc(&d_stats, 8 * sizeof(u64)));
    DRM_CUDA_TRY(cudaMemcpy(d_levels, ix.levels.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice));
    DRM_CUDA_TRY(cudaMemcpy(d_order, order.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice));
    DRM_CUDA_TRY(cudaMemcpy(d_upb, up_base.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice));
    if (n_up) {
        DRM_CUDA_TRY(cudaMemcpy(d_rv, row_vertex.data(), (size_t)n_up * sizeof(int), cudaMemcpyHostToDevice));
        DRM_CUDA_TRY(cudaMemcpy(d_rl, row_level.data(), (size_t)n_up * sizeof(int), cudaMemcpyHostToDevice));
    }
    DRM_CUDA_TRY(cudaMemset(d_locks, 0, (size_t)n * sizeof(int)));
    DRM_CUDA_TRY(cudaMemset(d_ctr, 0, (batches.size() + 1) * 4 * sizeof(unsigned)));
    DRM_CUDA_TRY(cudaMemset(d_stats, 0, 8 * sizeof(u64)));
    {
        const long long w0 = (long long)n * (p.row0_bytes / 4), wu = (long long)n_up * (p.rowu_bytes / 4);
        init_rows_kernel<<<(unsigned)((w0 + 255) / 256), 256>>>(p.rows0, n, p.row0_bytes, W0);
        if (n_up) init_rows_kernel<<<(unsigned)((wu + 255) / 256), 256>>>(p.up, n_up, p.rowu_bytes, Wu);
        sdc_kernel<<<(unsigned)((M_pq * 65536 + 255) / 256), 256>>>(d_cent, M_pq, dsub, d_sdc);
        DRM_CUDA_TRY(cudaGetLastError());
    }
    p.x = dx; p.order = d_order; p.centroids = d_cent; p.codes = d_codes; p.levels = d_levels; p.up_base = d_upb;
    p.sdc = d_sdc; p.locks = d_locks; p.requests = d_req; p.stats = d_stats;
    p.d = d; p.M = M_pq; p.dsub = dsub; p.efc = efc; p.cap_alloc = (efc + 1) & ~1;
    p.entry = entry; p.max_level = max_level;

    {
        auto insert_k = CS == 8 ? build_insert_kernel<8> : build_insert_kernel<16>;
        auto back_k = CS == 8 ? build_backlink_kernel<8> : build_backlink_kernel<16>;
        const size_t smem = (size_t)BUILD_WARPS * ((size_t)M_pq * 256 * sizeof(float) + (size_t)p.cap_alloc * sizeof(u64));
        if (smem > (size_t)prop.sharedMemPerBlockOptin) return fail(DRM_EUNSUPPORTED, "EFC / M_pq too large for the GPU builder's shared memory");
        DRM_CUDA_TRY(cudaFuncSetAttribute(insert_k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ_i = 0, occ_b = 0;
        DRM_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_i, insert_k, BUILD_WARPS * 32, smem));
        DRM_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_b, back_k, BUILD_WARPS * 32, 0));
        if (occ_i < 1 || occ_b < 1) return fail(DRM_ECUDA, "the GPU builder's kernels do not fit on this device");
        const long long full_i = (long long)occ_i * prop.multiProcessorCount, full_b = (long long)occ_b * prop.multiProcessorCount;
        cudaStream_t stream = nullptr;  // default stream: the kernels of consecutive batches must run in order
        for (size_t b = 0; b < batches.size(); ++b) {
            p.b0 = batches[b].first; p.b1 = batches[b].second;
            p.ctr = d_ctr + 4 * b;
            const long long nb = p.b1 - p.b0;
            const unsigned gi = (unsigned)std::min<long long>((nb + BUILD_WARPS - 1) / BUILD_WARPS, full_i);
            // at most W0 + levels * Wu requests per vertex; a persistent grid walks however many there are
            const unsigned gb = (unsigned)std::min<long long>((nb * W0 + BUILD_WARPS - 1) / BUILD_WARPS, full_b);
            insert_k<<<gi, BUILD_WARPS * 32, smem, stream>>>(p);
            back_k<<<gb, BUILD_WARPS * 32, 0, stream>>>(p);
        }
        DRM_CUDA_TRY(cudaGetLastError());
        DRM_CUDA_TRY(cudaDeviceSynchronize());
    }
    lap(3);

    // ---- FAISS arrays
    DRM_CUDA_TRY(mem.alloc(&d_off, (size_t)(n + 1) * sizeof(unsigned long long)));
    DRM_CUDA_TRY(mem.alloc(&d_nb, ix.neighbors.size() * sizeof(int)));
    DRM_CUDA_TRY(cudaMemcpy(d_off, ix.offsets.data(), (size_t)(n + 1) * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    {
        const long long t0 = (long long)n * W0, tu = (long long)n_up * Wu;
        export_rows0_kernel<<<(unsigned)((t0 + 255) / 256), 256>>>(p.rows0, n, p.row0_bytes, W0, d_off, d_nb);
        if (n_up) export_up_kernel<<<(unsigned)((tu + 255) / 256), 256>>>(p.up, n_up, p.rowu_bytes, Wu, W0, d_rv, d_rl, d_off, d_nb);
        DRM_CUDA_TRY(cudaGetLastError());
    }
    DRM_CUDA_TRY(cudaMemcpy(ix.neighbors.data(), d_nb, ix.neighbors.size() * sizeof(int), cudaMemcpyDeviceToHost));
    ix.entry_point = entry;
    ix.max_level = max_level;
    lap(4);
    if (verbose) {
        u64 st[8];
        cudaMemcpy(st, d_stats, sizeof st, cudaMemcpyDeviceToHost);
        fprintf(stderr, "[drm build_gpu] n=%lld batches=%zu upload %.2fs, PQ train %.2fs, encode %.2fs, graph %.2fs, export %.2fs; "
                        "hops/vertex %.1f, ADC evals/vertex %.1f, links appended %llu, rows re-shrunk %llu, lock spins %llu\n",
                (long long)n, batches.size(), tim[0], tim[1], tim[2], tim[3], tim[4], (double)st[0] / (double)n, (double)st[1] / (double)n,
                st[2], st[3], st[4]);
    }
    if (timing_s) memcpy(timing_s, tim, sizeof tim);
    return DRM_OK;
}

}  // namespace drm

extern "C" int drm_build_index_gpu(const float *x, int64_t n, int d, const float *train, int64_t n_train, int M_pq, int nbits,
                                   int M_hnsw, int efc, int device, uint32_t seed, const char *out_path, double *timing_s)
{
    if (!x || !out_path) return drm::fail(DRM_EINVAL, "null argument to drm_build_index_gpu");
    drm::HostIndex ix;
    int rc = drm::build_index_gpu(x, false, n, d, train, n_train, 1, M_pq, nbits, M_hnsw, efc, seed, device, ix, timing_s);
    if (rc != DRM_OK) return rc;
    return drm::write_index_file(out_path, ix);
}

extern "C" int drm_build_index_gpu_device(const float *x_dev, int64_t n, int d, int64_t train_step, int M_pq, int nbits, int M_hnsw,
                                          int efc, int device, uint32_t seed, const char *out_path, double *timing_s)
{
    if (!x_dev || !out_path) return drm::fail(DRM_EINVAL, "null argument to drm_build_index_gpu_device");
    drm::HostIndex ix;
    int rc = drm::build_index_gpu(x_dev, true, n, d, nullptr, 0, train_step, M_pq, nbits, M_hnsw, efc, seed, device, ix, timing_s);
    if (rc != DRM_OK) return rc;
    return drm::write_index_file(out_path, ix);
}
