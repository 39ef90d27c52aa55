// HNSW-PQ search on sm_100a: one warp per query.
//
// Replaces faiss::IndexHNSWPQ::search as called by the reference's faiss_search
// (src/hnswpq/search.cpp:13,39-40).  Algorithm = FAISS HNSW::search (greedy descent on the upper
// levels + ef-bounded level-0 search with the MinimaxHeap stop rule) under the CANONICAL tie rule of
// oracle/hnswpq_oracle.cpp (orc_search tie_rule=1): candidates and results are ordered by the total
// order (distance, id).  Under that order
//   * the candidate MinimaxHeap is "the cap = max(ef,k) smallest evaluated vertices", kept here as
//     a SORTED POOL of 64-bit keys  (float_bits(dis) << 32 | id << 1 | expanded)  in shared memory;
//   * the result heap is the first k pool entries;
//   * FAISS's VisitedTable is redundant: a vertex seen before is either still in the pool (found by
//     the same binary search that gives its insert position) or was rejected/evicted with a key
//     above the pool maximum, which only decreases - so it is rejected again.  No per-query visited
//     structure exists, which is what lets 24 queries share one SM.
//
// Memory: one hop = ONE contiguous row read (neighbour ids + the PQ codes of those neighbours,
// fused at index-open time, see index_image.cu); the ADC table (M x 256 fp32) lives in shared
// memory.  All fp32 arithmetic uses explicit round-to-nearest intrinsics in the oracle's order, so
// distances are bit-identical to the oracle's.
#include "common.h"

#include <atomic>
#include <cfloat>
#include <cstdlib>
#include <cstring>
#include <mutex>

#ifdef DRM_EMU
#define DRM_PREFETCH_L2(ptr) ((void)(ptr))
#else
#define DRM_PREFETCH_L2(ptr) asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr))
#endif

namespace drm {

namespace {

constexpr int WARPS_PER_CTA = 4;
constexpr unsigned FULL = 0xffffffffu;
typedef unsigned long long u64;

struct SearchParams {
    const float *q;
    long long n;
    int d, M, dsub;
    const float *centroids;
    const float4 *centroids_t;  // [M][dsub/4][256] float4, nullptr when dsub is not a multiple of 4
    const uint8_t *rows0;
    const uint8_t *up;
    int row0_bytes, rowu_bytes, W0, Wu, c0_off, cu_off;
    long long entry_vertex, entry_row, ntotal;
    int max_level;
    int k, ef, cap, cap_alloc;
    int prefetch_depth;   // rows L2-prefetched beyond the register read-ahead (0..2)
    float *D;
    long long *I;
    u64 *counters;        // nhops0, nhops_up, ndis_eval (nullable)
    unsigned *next_query; // work counter, zeroed before launch
    uint8_t entry_code[64];
};

template <int CS> struct Code;
template <> struct Code<8> { uint2 v; };
template <> struct Code<16> { uint4 v; };

template <int CS> __device__ __forceinline__ Code<CS> load_code(const uint8_t *p);
template <> __device__ __forceinline__ Code<8> load_code<8>(const uint8_t *p)
{
    Code<8> c; c.v = __ldg(reinterpret_cast<const uint2 *>(p)); return c;
}
template <> __device__ __forceinline__ Code<16> load_code<16>(const uint8_t *p)
{
    Code<16> c; c.v = __ldg(reinterpret_cast<const uint4 *>(p)); return c;
}

// dis = sum_m T[m][code[m]], m ascending, sequential fp32 adds (oracle adc_dis).
__device__ __forceinline__ float adc_word(const float *T, unsigned w, int m0, float acc, bool first)
{
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        float t = T[(m0 + b) * 256 + ((w >> (8 * b)) & 0xffu)];
        acc = (first && b == 0) ? t : __fadd_rn(acc, t);
    }
    return acc;
}
__device__ __forceinline__ float adc(const float *T, const Code<8> &c)
{
    float a = adc_word(T, c.v.x, 0, 0.f, true);
    return adc_word(T, c.v.y, 4, a, false);
}
__device__ __forceinline__ float adc(const float *T, const Code<16> &c)
{
    float a = adc_word(T, c.v.x, 0, 0.f, true);
    a = adc_word(T, c.v.y, 4, a, false);
    a = adc_word(T, c.v.z, 8, a, false);
    return adc_word(T, c.v.w, 12, a, false);
}
template <int CS> __device__ __forceinline__ float adc_bytes(const float *T, const uint8_t *code)
{
    float a = T[code[0]];
    for (int m = 1; m < CS; ++m) a = __fadd_rn(a, T[m * 256 + code[m]]);
    return a;
}

__device__ __forceinline__ u64 make_key(float d, int id) { return ((u64)__float_as_uint(d) << 32) | ((u64)(unsigned)id << 1); }

// T[m][j] = sum_t (x[m*dsub+t] - c[m][j][t])^2, t ascending, one fma per term (oracle adc_table).
__device__ __forceinline__ void build_lut(float *T, const float *__restrict__ x, const float *__restrict__ cent, int M, int dsub, int lane)
{
    for (int m = 0; m < M; ++m) {
        const float *xm = x + m * dsub;
        for (int j = lane; j < 256; j += 32) {
            const float *c = cent + ((size_t)(m * 256 + j)) * dsub;
            float acc = 0.f;
            if ((dsub & 3) == 0) {
                for (int t = 0; t < dsub; t += 4) {
                    float4 cv = __ldg(reinterpret_cast<const float4 *>(c + t));
                    float4 xv = __ldg(reinterpret_cast<const float4 *>(xm + t));
                    float e0 = __fsub_rn(xv.x, cv.x); acc = __fmaf_rn(e0, e0, acc);
                    float e1 = __fsub_rn(xv.y, cv.y); acc = __fmaf_rn(e1, e1, acc);
                    float e2 = __fsub_rn(xv.z, cv.z); acc = __fmaf_rn(e2, e2, acc);
                    float e3 = __fsub_rn(xv.w, cv.w); acc = __fmaf_rn(e3, e3, acc);
                }
            } else {
                for (int t = 0; t < dsub; ++t) {
                    float e = __fsub_rn(__ldg(xm + t), __ldg(c + t));
                    acc = __fmaf_rn(e, e, acc);
                }
            }
            T[m * 256 + j] = acc;
        }
    }
}

__device__ __forceinline__ u64 warp_min_u64(u64 v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        u64 other = __shfl_xor_sync(FULL, v, o);
        v = other < v ? other : v;
    }
    return v;
}

// NPL = neighbours per lane at level 0 (W0 <= 32*NPL).
template <int CS, int NPL>
__global__ void __launch_bounds__(WARPS_PER_CTA * 32) hnswpq_search_kernel(const SearchParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t per_warp = (size_t)p.M * 256 * sizeof(float) + (size_t)p.cap_alloc * sizeof(u64);
    float *T = reinterpret_cast<float *>(smem_raw + warp * per_warp);
    u64 *pool = reinterpret_cast<u64 *>(smem_raw + warp * per_warp + (size_t)p.M * 256 * sizeof(float));
    const int cap = p.cap;

    for (;;) {
        unsigned qi = 0;
        if (lane == 0) qi = atomicAdd(p.next_query, 1u);
        qi = __shfl_sync(FULL, qi, 0);
        if ((long long)qi >= p.n) break;

        __syncwarp();
        build_lut(T, p.q + (size_t)qi * p.d, p.centroids, p.M, p.dsub, lane);
        __syncwarp();

        unsigned nhops0 = 0, nhops_up = 0, ndis = 0;

        // ---------------- greedy descent, levels max_level .. 1 (FAISS greedy_update_nearest)
        float d_near = adc_bytes<CS>(T, p.entry_code);
        long long vertex = p.entry_vertex;
        if (p.max_level >= 1) {
            long long r = p.entry_row;
            for (int level = p.max_level; level >= 1; --level) {
                for (;;) {
                    const uint8_t *row = p.up + (size_t)r * p.rowu_bytes;
                    int nr = -1;
                    float dj = __uint_as_float(0x7f800000u);
                    if (lane < p.Wu) nr = __ldg(reinterpret_cast<const int *>(row + 16) + lane);
                    unsigned neg = __ballot_sync(FULL, lane < p.Wu && nr < 0);
                    int first_neg = neg ? (__ffs(neg) - 1) : p.Wu;
                    if (lane < first_neg) dj = adc(T, load_code<CS>(row + p.cu_off + CS * lane));
                    u64 best = warp_min_u64(((u64)__float_as_uint(dj) << 32) | (unsigned)lane);
                    float dmin = __uint_as_float((unsigned)(best >> 32));
                    ++nhops_up;
                    if (dmin < d_near) {
                        d_near = dmin;
                        r = __shfl_sync(FULL, nr, (int)(best & 31u));
                    } else break;
                }
                if (level > 1) r -= 1;  // same vertex, one level down (rows of a vertex are contiguous)
            }
            vertex = __ldg(reinterpret_cast<const int *>(p.up + (size_t)r * p.rowu_bytes));
        }

        // ---------------- level 0: sorted-pool search
        int size = 1, cur = 0;
        if (lane == 0) pool[0] = make_key(d_near, (int)vertex);
        __syncwarp();
        while (cur >= 0) {
            const u64 k0 = pool[cur];
            const unsigned v0 = (unsigned)(k0 & 0xffffffffu) >> 1;
            if (cap > p.ef) {
                // MinimaxHeap::count_below(d0) >= efSearch (only reachable when k > ef)
                const unsigned d0 = (unsigned)(k0 >> 32);
                int below = 0;
                for (int b = 0; b < size; b += 32) {
                    int i = b + lane;
                    below += __popc(__ballot_sync(FULL, i < size && (unsigned)(pool[i] >> 32) < d0));
                }
                if (below >= p.ef) break;
            }
            const uint8_t *row = p.rows0 + (size_t)v0 * p.row0_bytes;
            int nid[NPL];
            Code<CS> code[NPL];
#pragma unroll
            for (int c = 0; c < NPL; ++c) {
                int slot = lane + 32 * c;
                nid[c] = slot < p.W0 ? __ldg(reinterpret_cast<const int *>(row) + slot) : -1;
                if (slot < p.W0) code[c] = load_code<CS>(row + p.c0_off + CS * slot);
            }
            __syncwarp();
            if (lane == 0) pool[cur] = k0 | 1ull;  // expanded
            // speculative L2 prefetch of the likeliest next row (first unexpanded entry after cur)
            {
                int i = cur + 1 + lane;
                u64 kk = i < size ? pool[i] : 1ull;
                unsigned m = __ballot_sync(FULL, !(kk & 1ull));
                if (m) {
                    unsigned nv = (unsigned)(__shfl_sync(FULL, kk, __ffs(m) - 1) & 0xffffffffu) >> 1;
                    const uint8_t *nrow = p.rows0 + (size_t)nv * p.row0_bytes;
                    if (lane * 128 < p.row0_bytes) DRM_PREFETCH_L2(nrow + lane * 128);
                }
            }
            __syncwarp();
            ++nhops0;

            u64 key[NPL];
            int lo[NPL];
            bool cand[NPL];
            bool row_open = true;
#pragma unroll
            for (int c = 0; c < NPL; ++c) {
                unsigned neg = __ballot_sync(FULL, nid[c] < 0);
                int first_neg = neg ? (__ffs(neg) - 1) : 32;
                bool valid = row_open && lane < first_neg;
                if (neg) row_open = false;
                ndis += __popc(__ballot_sync(FULL, valid));
                key[c] = 0;
                if (valid) key[c] = make_key(adc(T, code[c]), nid[c]);
                cand[c] = valid;
            }
            const bool full = size == cap;
            const u64 maxkey = full ? (pool[cap - 1] & ~1ull) : ~0ull;
            const int iters = 32 - __clz(size);
#pragma unroll
            for (int c = 0; c < NPL; ++c) {
                cand[c] = cand[c] && key[c] < maxkey;
                int l = 0, h = size;
                for (int it = 0; it < iters; ++it) {
                    if (l < h) {
                        int mid = (l + h) >> 1;
                        if ((pool[mid] & ~1ull) < key[c]) l = mid + 1; else h = mid;
                    }
                }
                lo[c] = l;
                if (cand[c] && l < size && (pool[l] & ~1ull) == key[c]) cand[c] = false;  // already in the pool
            }
            int hint = cur + 1;
#pragma unroll
            for (int c = 0; c < NPL; ++c) {
                unsigned mask = __ballot_sync(FULL, cand[c]);
                while (mask) {
                    const int src = __ffs(mask) - 1;
                    mask &= mask - 1;
                    const u64 kin = __shfl_sync(FULL, key[c], src);
                    const int pin = __shfl_sync(FULL, lo[c], src);
                    if (pin >= cap) continue;
                    if (pin < size && (pool[pin] & ~1ull) == kin) continue;  // duplicate id inside one row
                    const int nsz = size < cap ? size + 1 : cap;
                    for (int top = nsz; top > pin + 1; top -= 32) {
                        int i = top - 1 - lane;
                        u64 tmp = 0;
                        if (i > pin) tmp = pool[i - 1];
                        __syncwarp();
                        if (i > pin) pool[i] = tmp;
                        __syncwarp();
                    }
                    __syncwarp();  // every lane has finished its duplicate test on pool[pin] (lanes are not in lockstep)
                    if (lane == 0) pool[pin] = kin;
                    __syncwarp();
                    size = nsz;
#pragma unroll
                    for (int c2 = 0; c2 < NPL; ++c2)
                        if (key[c2] > kin) lo[c2] += 1;
                    hint = pin < hint ? pin : hint;
                }
            }
            cur = -1;
            for (int b = hint; b < size; b += 32) {
                int i = b + lane;
                unsigned m = __ballot_sync(FULL, i < size && !(pool[i] & 1ull));
                if (m) { cur = b + __ffs(m) - 1; break; }
            }
        }

        // ---------------- results: the first k pool entries, ascending (distance, id)
        float *Dq = p.D + (size_t)qi * p.k;
        long long *Iq = p.I + (size_t)qi * p.k;
        for (int i = lane; i < p.k; i += 32) {
            if (i < size) {
                u64 kk = pool[i];
                Dq[i] = __uint_as_float((unsigned)(kk >> 32));
                Iq[i] = (long long)((unsigned)(kk & 0xffffffffu) >> 1);
            } else {
                Dq[i] = FLT_MAX;
                Iq[i] = -1;
            }
        }
        if (p.counters && lane == 0) {
            atomicAdd(p.counters + 0, (u64)nhops0);
            atomicAdd(p.counters + 1, (u64)nhops_up);
            atomicAdd(p.counters + 2, (u64)ndis);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Register-pool kernel: the same algorithm with the sorted pool held in REGISTERS, NS keys per lane
// (pool position i lives in lane i / NS, slot i % NS; cap <= 32 * NS).  Inserting a key is a rank
// computation (per-lane compares + one REDUX add) and a one-position shift (selects + one shuffle for the
// lane boundary) - no shared-memory traffic, no loops over the pool, no __syncwarp.  It also reads the
// row of the likeliest next vertex (the first unexpanded pool entry) before it processes the current row, so the DRAM
// latency of the next hop overlaps the ADC evaluations and insertions of this one; when an insertion
// overtakes that entry the row is simply fetched again later (by then it is an L2 hit).
// Preconditions (launch_search falls back to the shared-memory kernel otherwise): W0 <= 32, k <= ef <= 256.
constexpr u64 POOL_INF = ~0ull;

template <int NS> __device__ __forceinline__ bool pool_first_unexpanded(const u64 (&k)[NS], int lane, int &L, u64 &key, u64 &mine, unsigned &bal)
{
    mine = POOL_INF;
#pragma unroll
    for (int s = NS - 1; s >= 0; --s)
        if (!(k[s] & 1ull)) mine = k[s];
    bal = __ballot_sync(FULL, !(mine & 1ull));
    if (!bal) return false;
    L = __ffs(bal) - 1;
    key = __shfl_sync(FULL, mine, L);
    return true;
}
template <int NS> __device__ __forceinline__ bool pool_first_unexpanded(const u64 (&k)[NS], int lane, int &L, u64 &key)
{
    u64 mine; unsigned bal;
    return pool_first_unexpanded<NS>(k, lane, L, key, mine, bal);
}

__device__ __forceinline__ void build_lut_t(float *T, const float *__restrict__ x, const float4 *__restrict__ cent_t, int M, int dsub, int lane)
{
    const int q4 = dsub >> 2;
    for (int m = 0; m < M; ++m) {
        const float *xm = x + m * dsub;
        for (int j = lane; j < 256; j += 32) {
            float acc = 0.f;
            for (int t4 = 0; t4 < q4; ++t4) {
                const float4 cv = __ldg(cent_t + ((size_t)(m * q4 + t4) * 256 + j));
                const float4 xv = __ldg(reinterpret_cast<const float4 *>(xm) + t4);
                float e0 = __fsub_rn(xv.x, cv.x); acc = __fmaf_rn(e0, e0, acc);
                float e1 = __fsub_rn(xv.y, cv.y); acc = __fmaf_rn(e1, e1, acc);
                float e2 = __fsub_rn(xv.z, cv.z); acc = __fmaf_rn(e2, e2, acc);
                float e3 = __fsub_rn(xv.w, cv.w); acc = __fmaf_rn(e3, e3, acc);
            }
            T[m * 256 + j] = acc;
        }
    }
}

template <int NS, bool FULLCAP>
__device__ __forceinline__ void pool_insert_mask(u64 (&k)[NS], u64 &maxkey, const u64 key, unsigned mask, int lane, int base, int cap,
                                                 int max_lane, int max_slot)
{
    while (mask) {
        const int src = __ffs(mask) - 1;
        mask &= mask - 1;
        const u64 kin = __shfl_sync(FULL, key, src);
        if (kin >= maxkey) continue;
        // Sorted insert without a rank computation.  kin has flag 0 and pool keys are (even base | flag), so `k[s] < kin`
        // already ignores the flag; the pool is sorted, so lt[] is a prefix of the positions and
        //   new[i] = lt[i] ? old[i] : (lt[i-1] ? kin : old[i-1])
        // with position i-1 living in the previous lane for slot 0 (one 64-bit shuffle).  Duplicate test on the low word
        // only: (id << 1 | flag) identifies the vertex, and a vertex has one distance.
        const unsigned id1 = (unsigned)kin | 1u;
        bool lt[NS];
        bool dup = false;
#pragma unroll
        for (int s = 0; s < NS; ++s) {
            lt[s] = k[s] < kin;
            dup = dup || ((unsigned)k[s] | 1u) == id1;
        }
        if (__any_sync(FULL, dup)) continue;  // already in the pool (or the same id twice in one row)
        const u64 pk = __shfl_up_sync(FULL, k[NS - 1], 1);
        const bool plt = lane == 0 || pk < kin;
#pragma unroll
        for (int s = NS - 1; s >= 1; --s) k[s] = lt[s] ? k[s] : (lt[s - 1] ? kin : k[s - 1]);
        k[0] = lt[0] ? k[0] : (plt ? kin : pk);
        if (FULLCAP) {
            maxkey = __shfl_sync(FULL, k[NS - 1], 31) & ~1ull;
        } else {
#pragma unroll
            for (int s = 0; s < NS; ++s)
                if (base + s >= cap) k[s] = POOL_INF;
            u64 lastk = k[0];
#pragma unroll
            for (int s = 1; s < NS; ++s)
                if (s == max_slot) lastk = k[s];
            maxkey = __shfl_sync(FULL, lastk, max_lane) & ~1ull;
        }
    }
}

template <int NS> __device__ __forceinline__ void pool_mark_first_unexpanded(u64 (&k)[NS], int lane, int L)
{
    bool done = lane != L;
#pragma unroll
    for (int s = 0; s < NS; ++s) {
        const bool hit = !done && !(k[s] & 1ull);
        if (hit) k[s] |= 1ull;
        done = done || hit;
    }
}

// FULLCAP: cap == 32 * NS (the pool fills the registers exactly: no tail to keep at POOL_INF, the maximum is lane 31's last slot).
// VIS: a per-warp direct-mapped cache of evaluated vertex ids in shared memory (VIS_SLOTS 16-bit tags: slot = id % VIS_SLOTS,
// tag = id / VIS_SLOTS, exact while ntotal <= VIS_SLOTS * 65535).  A hit means "evaluated before", and a vertex evaluated
// before is either still in the pool or was rejected / evicted above the pool maximum (which only decreases), so it can be
// skipped without changing the result; a miss after an eviction from the cache just re-evaluates, which the duplicate test
// and the pool maximum handle as before.  It removes the repeat ADC evaluations and insert attempts (about a quarter).
// VIS = number of slots (0 = no cache): 1024 slots leave room for 5 CTAs per SM, 512 for 6.
template <int CS, int NS, bool FULLCAP, int VIS>
__global__ void __launch_bounds__(WARPS_PER_CTA * 32, VIS > 512 ? 5 : 6) hnswpq_search_reg_kernel(const SearchParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *T = reinterpret_cast<float *>(smem_raw) + (size_t)warp * p.M * 256;
    constexpr int VIS_SLOTS = VIS > 0 ? VIS : 1;
    unsigned short *vis = reinterpret_cast<unsigned short *>(smem_raw + (size_t)WARPS_PER_CTA * p.M * 256 * sizeof(float)) + (size_t)warp * VIS_SLOTS;
    const int cap = p.cap;
    const int base = lane * NS;
    const int max_lane = (cap - 1) / NS, max_slot = (cap - 1) % NS;

    for (;;) {
        unsigned qi = 0;
        if (lane == 0) qi = atomicAdd(p.next_query, 1u);
        qi = __shfl_sync(FULL, qi, 0);
        if ((long long)qi >= p.n) break;

        __syncwarp();
        if (p.centroids_t) build_lut_t(T, p.q + (size_t)qi * p.d, p.centroids_t, p.M, p.dsub, lane);
        else build_lut(T, p.q + (size_t)qi * p.d, p.centroids, p.M, p.dsub, lane);
        if (VIS) {
            uint4 *v4 = reinterpret_cast<uint4 *>(vis);
#pragma unroll
            for (int i = 0; i < (VIS_SLOTS * 2 / 16 + 31) / 32; ++i) v4[lane + 32 * i] = make_uint4(~0u, ~0u, ~0u, ~0u);
        }
        __syncwarp();

        unsigned nhops0 = 0, nhops_up = 0, ndis = 0;

        // ---------------- greedy descent, levels max_level .. 1 (FAISS greedy_update_nearest)
        float d_near = adc_bytes<CS>(T, p.entry_code);
        long long vertex = p.entry_vertex;
        if (p.max_level >= 1) {
            long long r = p.entry_row;
            for (int level = p.max_level; level >= 1; --level) {
                for (;;) {
                    const uint8_t *row = p.up + (size_t)r * p.rowu_bytes;
                    int nr = -1;
                    float dj = __uint_as_float(0x7f800000u);
                    if (lane < p.Wu) nr = __ldg(reinterpret_cast<const int *>(row + 16) + lane);
                    unsigned neg = __ballot_sync(FULL, lane < p.Wu && nr < 0);
                    int first_neg = neg ? (__ffs(neg) - 1) : p.Wu;
                    if (lane < first_neg) dj = adc(T, load_code<CS>(row + p.cu_off + CS * lane));
                    u64 best = warp_min_u64(((u64)__float_as_uint(dj) << 32) | (unsigned)lane);
                    float dmin = __uint_as_float((unsigned)(best >> 32));
                    ++nhops_up;
                    if (dmin < d_near) {
                        d_near = dmin;
                        r = __shfl_sync(FULL, nr, (int)(best & 31u));
                    } else break;
                }
                if (level > 1) r -= 1;
            }
            vertex = __ldg(reinterpret_cast<const int *>(p.up + (size_t)r * p.rowu_bytes));
        }

        // ---------------- level 0
        if (VIS && lane == 0) vis[(unsigned)vertex % VIS_SLOTS] = (unsigned short)((unsigned)vertex / VIS_SLOTS);
        u64 k[NS];
#pragma unroll
        for (int s = 0; s < NS; ++s) k[s] = POOL_INF;
        if (lane == 0) k[0] = make_key(d_near, (int)vertex);
        u64 maxkey = POOL_INF & ~1ull;
        unsigned spec_v = 0xffffffffu;
        int nid_s = -1;
        Code<CS> code_s;
        code_s.v = {};
        for (;;) {
            int L; u64 k0;
            if (!pool_first_unexpanded<NS>(k, lane, L, k0)) break;
            const unsigned v0 = (unsigned)(k0 & 0xffffffffu) >> 1;
            pool_mark_first_unexpanded<NS>(k, lane, L);
            int nid; Code<CS> code;
            if (v0 == spec_v) { nid = nid_s; code = code_s; }
            else {
                const uint8_t *row = p.rows0 + (size_t)v0 * p.row0_bytes;
                nid = lane < p.W0 ? __ldg(reinterpret_cast<const int *>(row) + lane) : -1;
                code.v = {};
                if (lane < p.W0) code = load_code<CS>(row + p.c0_off + CS * lane);
            }
            {   // read ahead: the row of the next entry in line
                int L2; u64 k2, mine; unsigned bal;
                if (pool_first_unexpanded<NS>(k, lane, L2, k2, mine, bal)) {
                    spec_v = (unsigned)(k2 & 0xffffffffu) >> 1;
                    const uint8_t *nrow = p.rows0 + (size_t)spec_v * p.row0_bytes;
                    nid_s = lane < p.W0 ? __ldg(reinterpret_cast<const int *>(nrow) + lane) : -1;
                    if (lane < p.W0) code_s = load_code<CS>(nrow + p.c0_off + CS * lane);
                    if (p.prefetch_depth > 0) {
                        // L2 prefetch of the rows further down the line (the first unexpanded entry of the next lanes that
                        // have one): HBM has the headroom (8 % used), and a row that an insertion pushes back is still
                        // expanded later, so the lines are rarely wasted.
                        unsigned rest = bal & (bal - 1);  // drop lane L2
                        const int la = rest ? __ffs(rest) - 1 : -1;
                        rest &= rest - 1;
                        const int lb = (rest && p.prefetch_depth > 1) ? __ffs(rest) - 1 : -1;
                        if (lane == la || lane == lb) {
                            const uint8_t *prow = p.rows0 + (size_t)((unsigned)(mine & 0xffffffffu) >> 1) * p.row0_bytes;
                            for (int o = 0; o < p.row0_bytes; o += 128) DRM_PREFETCH_L2(prow + o);
                        }
                    }
                } else spec_v = 0xffffffffu;
            }
            ++nhops0;

            const unsigned neg = __ballot_sync(FULL, nid < 0);
            const int first_neg = neg ? (__ffs(neg) - 1) : 32;
            bool valid = lane < first_neg;
            if (VIS) {
                if (valid) {
                    const unsigned slot = (unsigned)nid % VIS_SLOTS;
                    const unsigned short tag = (unsigned short)((unsigned)nid / VIS_SLOTS);
                    if (vis[slot] == tag) valid = false;
                    else vis[slot] = tag;
                }
                const unsigned fresh = __ballot_sync(FULL, valid);
                ndis += (unsigned)__popc(fresh);
                if (!fresh) continue;
            } else {
                ndis += (unsigned)first_neg;
            }
            u64 key = 0;
            if (valid) key = make_key(adc(T, code), nid);
            const unsigned mask = __ballot_sync(FULL, valid && key < maxkey);
            pool_insert_mask<NS, FULLCAP>(k, maxkey, key, mask, lane, base, cap, max_lane, max_slot);
        }

        // ---------------- results: the first k pool entries, ascending (distance, id)
        float *Dq = p.D + (size_t)qi * p.k;
        long long *Iq = p.I + (size_t)qi * p.k;
#pragma unroll
        for (int s = 0; s < NS; ++s) {
            const int i = base + s;
            if (i < p.k) {
                const u64 kk = k[s];
                const bool filled = kk != POOL_INF;
                Dq[i] = filled ? __uint_as_float((unsigned)(kk >> 32)) : FLT_MAX;
                Iq[i] = filled ? (long long)((unsigned)(kk & 0xffffffffu) >> 1) : -1ll;
            }
        }
        // positions beyond the register pool (k > 32 * NS cannot happen: k <= cap <= 32 * NS)
        if (p.counters && lane == 0) {
            atomicAdd(p.counters + 0, (u64)nhops0);
            atomicAdd(p.counters + 1, (u64)nhops_up);
            atomicAdd(p.counters + 2, (u64)ndis);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// PAIR variant of the register-pool kernel (DRM_SEARCH_KERNEL=pair): every iteration evaluates the rows of the first TWO
// unexpanded pool entries at once - lanes 0..15 take the first 16 slots of the first row, lanes 16..31 those of the second -
// then commits the first row's candidates and, if the second entry is still the first unexpanded one afterwards (56 % of the
// hops on the bench workload, oracle/hnswpq_oracle.cpp: orc_spec2_counters), the second row's as well.  The sequence of pool
// states is exactly the sequential one: a speculative row that is not next is dropped (it touched neither the pool nor the
// visited cache), and one that is next sees the pool, the maximum and the duplicate test as they are after the first commit.
// A first row with more than 16 live neighbours is re-read by all 32 lanes (no speculation in that iteration); a second row
// with more than 16 is not committed.  Rows further down the queue are prefetched into L2.
template <int CS, int NS, bool FULLCAP, int VIS>
__global__ void __launch_bounds__(WARPS_PER_CTA * 32, VIS > 512 ? 5 : 6) hnswpq_search_pair_kernel(const SearchParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *T = reinterpret_cast<float *>(smem_raw) + (size_t)warp * p.M * 256;
    constexpr int VIS_SLOTS = VIS > 0 ? VIS : 1;
    unsigned short *vis = reinterpret_cast<unsigned short *>(smem_raw + (size_t)WARPS_PER_CTA * p.M * 256 * sizeof(float)) + (size_t)warp * VIS_SLOTS;
    const int cap = p.cap;
    const int base = lane * NS;
    const int max_lane = (cap - 1) / NS, max_slot = (cap - 1) % NS;
    const int half = lane >> 4, hslot = lane & 15;

    for (;;) {
        unsigned qi = 0;
        if (lane == 0) qi = atomicAdd(p.next_query, 1u);
        qi = __shfl_sync(FULL, qi, 0);
        if ((long long)qi >= p.n) break;

        __syncwarp();
        if (p.centroids_t) build_lut_t(T, p.q + (size_t)qi * p.d, p.centroids_t, p.M, p.dsub, lane);
        else build_lut(T, p.q + (size_t)qi * p.d, p.centroids, p.M, p.dsub, lane);
        if (VIS) {
            uint4 *v4 = reinterpret_cast<uint4 *>(vis);
#pragma unroll
            for (int i = 0; i < (VIS_SLOTS * 2 / 16 + 31) / 32; ++i) v4[lane + 32 * i] = make_uint4(~0u, ~0u, ~0u, ~0u);
        }
        __syncwarp();

        unsigned nhops0 = 0, nhops_up = 0, ndis = 0;

        // ---------------- greedy descent, levels max_level .. 1 (as in the sequential kernel)
        float d_near = adc_bytes<CS>(T, p.entry_code);
        long long vertex = p.entry_vertex;
        if (p.max_level >= 1) {
            long long r = p.entry_row;
            for (int level = p.max_level; level >= 1; --level) {
                for (;;) {
                    const uint8_t *row = p.up + (size_t)r * p.rowu_bytes;
                    int nr = -1;
                    float dj = __uint_as_float(0x7f800000u);
                    if (lane < p.Wu) nr = __ldg(reinterpret_cast<const int *>(row + 16) + lane);
                    unsigned neg = __ballot_sync(FULL, lane < p.Wu && nr < 0);
                    int first_neg = neg ? (__ffs(neg) - 1) : p.Wu;
                    if (lane < first_neg) dj = adc(T, load_code<CS>(row + p.cu_off + CS * lane));
                    u64 best = warp_min_u64(((u64)__float_as_uint(dj) << 32) | (unsigned)lane);
                    float dmin = __uint_as_float((unsigned)(best >> 32));
                    ++nhops_up;
                    if (dmin < d_near) {
                        d_near = dmin;
                        r = __shfl_sync(FULL, nr, (int)(best & 31u));
                    } else break;
                }
                if (level > 1) r -= 1;
            }
            vertex = __ldg(reinterpret_cast<const int *>(p.up + (size_t)r * p.rowu_bytes));
        }

        // ---------------- level 0
        if (VIS && lane == 0) vis[(unsigned)vertex % VIS_SLOTS] = (unsigned short)((unsigned)vertex / VIS_SLOTS);
        u64 k[NS];
#pragma unroll
        for (int s = 0; s < NS; ++s) k[s] = POOL_INF;
        if (lane == 0) k[0] = make_key(d_near, (int)vertex);
        u64 maxkey = POOL_INF & ~1ull;
        for (;;) {
            // first and second unexpanded entry of every lane, then of the pool (positions are lane-major)
            u64 mine = POOL_INF, mine2 = POOL_INF;
#pragma unroll
            for (int s = NS - 1; s >= 0; --s)
                if (!(k[s] & 1ull)) { mine2 = mine; mine = k[s]; }
            const unsigned bal = __ballot_sync(FULL, !(mine & 1ull));
            if (!bal) break;
            const int L1 = __ffs(bal) - 1;
            const u64 k1 = __shfl_sync(FULL, mine, L1);
            const u64 s2 = __shfl_sync(FULL, mine2, L1);
            const unsigned rest = bal & (bal - 1);
            const int Lb = rest ? __ffs(rest) - 1 : L1;
            const u64 nb = __shfl_sync(FULL, mine, Lb);
            const bool same_lane = !(s2 & 1ull);
            const bool has2 = same_lane || rest != 0;
            const u64 k2 = same_lane ? s2 : nb;
            pool_mark_first_unexpanded<NS>(k, lane, L1);
            const unsigned v1 = (unsigned)(k1 & 0xffffffffu) >> 1, v2 = (unsigned)(k2 & 0xffffffffu) >> 1;

            // the two half rows
            const bool load = hslot < p.W0 && (half == 0 || has2);
            const uint8_t *row = p.rows0 + (size_t)(half ? v2 : v1) * p.row0_bytes;
            int nid = -1;
            Code<CS> code;
            code.v = {};
            if (load) {
                nid = __ldg(reinterpret_cast<const int *>(row) + hslot);
                code = load_code<CS>(row + p.c0_off + CS * hslot);
            }
            {   // L2 prefetch of the rows of the next entries in the queue (other lanes' first unexpanded entries)
                unsigned q = same_lane ? rest : (rest & (rest - 1));
                const int la = q ? __ffs(q) - 1 : -1;
                q &= q - 1;
                const int lb = q ? __ffs(q) - 1 : -1;
                if (lane == la || lane == lb) {
                    const uint8_t *prow = p.rows0 + (size_t)((unsigned)(mine & 0xffffffffu) >> 1) * p.row0_bytes;
                    for (int o = 0; o < p.row0_bytes; o += 128) DRM_PREFETCH_L2(prow + o);
                }
            }
            ++nhops0;
            const unsigned neg = __ballot_sync(FULL, nid < 0);
            const unsigned neg1 = neg & 0xffffu, neg2 = neg >> 16;
            const int fn1 = neg1 ? __ffs(neg1) - 1 : 16, fn2 = neg2 ? __ffs(neg2) - 1 : 16;

            if (fn1 == 16 && p.W0 > 16) {
                // the first row may have more than 16 live neighbours: all 32 lanes take it, nothing speculative this time
                const uint8_t *r1 = p.rows0 + (size_t)v1 * p.row0_bytes;
                nid = lane < p.W0 ? __ldg(reinterpret_cast<const int *>(r1) + lane) : -1;
                code.v = {};
                if (lane < p.W0) code = load_code<CS>(r1 + p.c0_off + CS * lane);
                const unsigned ng = __ballot_sync(FULL, nid < 0);
                const int first_neg = ng ? (__ffs(ng) - 1) : 32;
                bool valid = lane < first_neg;
                if (VIS) {
                    if (valid) {
                        const unsigned slot = (unsigned)nid % VIS_SLOTS;
                        const unsigned short tag = (unsigned short)((unsigned)nid / VIS_SLOTS);
                        if (vis[slot] == tag) valid = false;
                        else vis[slot] = tag;
                    }
                }
                const unsigned fresh = __ballot_sync(FULL, valid);
                ndis += (unsigned)__popc(fresh);
                u64 key = 0;
                if (valid) key = make_key(adc(T, code), nid);
                const unsigned mask = __ballot_sync(FULL, valid && key < maxkey);
                pool_insert_mask<NS, FULLCAP>(k, maxkey, key, mask, lane, base, cap, max_lane, max_slot);
                continue;
            }

            const bool spec_ok = has2 && !(fn2 == 16 && p.W0 > 16);
            bool valid = half == 0 ? hslot < fn1 : (spec_ok && hslot < fn2);
            unsigned slot = 0;
            unsigned short tag = 0;
            if (VIS) {
                if (valid) {
                    slot = (unsigned)nid % VIS_SLOTS;
                    tag = (unsigned short)((unsigned)nid / VIS_SLOTS);
                    if (vis[slot] == tag) valid = false;
                }
                __syncwarp();
                if (valid && half == 0) vis[slot] = tag;  // the speculative half marks nothing until it is committed
            }
            const unsigned fresh = __ballot_sync(FULL, valid);
            ndis += (unsigned)__popc(fresh & 0xffffu);
            u64 key = 0;
            if (valid) key = make_key(adc(T, code), nid);
            {
                const unsigned mask = __ballot_sync(FULL, valid && key < maxkey) & 0xffffu;
                pool_insert_mask<NS, FULLCAP>(k, maxkey, key, mask, lane, base, cap, max_lane, max_slot);
            }
            if (spec_ok) {
                int Lx; u64 kx;
                if (pool_first_unexpanded<NS>(k, lane, Lx, kx) && kx == k2) {
                    pool_mark_first_unexpanded<NS>(k, lane, Lx);
                    ++nhops0;
                    if (VIS && valid && half == 1) vis[slot] = tag;
                    ndis += (unsigned)__popc(fresh >> 16);
                    const unsigned mask = __ballot_sync(FULL, valid && key < maxkey) & 0xffff0000u;
                    pool_insert_mask<NS, FULLCAP>(k, maxkey, key, mask, lane, base, cap, max_lane, max_slot);
                }
            }
        }

        // ---------------- results: the first k pool entries, ascending (distance, id)
        float *Dq = p.D + (size_t)qi * p.k;
        long long *Iq = p.I + (size_t)qi * p.k;
#pragma unroll
        for (int s = 0; s < NS; ++s) {
            const int i = base + s;
            if (i < p.k) {
                const u64 kk = k[s];
                const bool filled = kk != POOL_INF;
                Dq[i] = filled ? __uint_as_float((unsigned)(kk >> 32)) : FLT_MAX;
                Iq[i] = filled ? (long long)((unsigned)(kk & 0xffffffffu) >> 1) : -1ll;
            }
        }
        if (p.counters && lane == 0) {
            atomicAdd(p.counters + 0, (u64)nhops0);
            atomicAdd(p.counters + 1, (u64)nhops_up);
            atomicAdd(p.counters + 2, (u64)ndis);
        }
    }
}

template <int CS, int NS, bool FULLCAP, int VIS>
int launch_pair_tfv(const SearchParams &p, size_t smem, int device, cudaStream_t stream)
{
    auto kern = hnswpq_search_pair_kernel<CS, NS, FULLCAP, VIS>;
    smem += (size_t)WARPS_PER_CTA * VIS * sizeof(unsigned short);
    DRM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0, sms = 0;
    DRM_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WARPS_PER_CTA * 32, smem));
    DRM_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    if (per_sm < 1) return fail(DRM_EUNSUPPORTED, "search kernel does not fit on an SM with these parameters");
    long long want = (p.n + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
    long long grid = (long long)per_sm * sms;
    if (want < grid) grid = want;
    kern<<<(unsigned)grid, WARPS_PER_CTA * 32, smem, stream>>>(p);
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

template <int CS, int NS>
int launch_pair_t(const SearchParams &p, size_t smem, int device, cudaStream_t stream)
{
    const bool vis = p.ntotal < 512ll * 65535;
    if (p.cap == 32 * NS) return vis ? launch_pair_tfv<CS, NS, true, 512>(p, smem, device, stream) : launch_pair_tfv<CS, NS, true, 0>(p, smem, device, stream);
    return vis ? launch_pair_tfv<CS, NS, false, 512>(p, smem, device, stream) : launch_pair_tfv<CS, NS, false, 0>(p, smem, device, stream);
}

template <int CS, int NS, bool FULLCAP, int VIS>
int launch_reg_tfv(const SearchParams &p, size_t smem, int device, cudaStream_t stream)
{
    auto kern = hnswpq_search_reg_kernel<CS, NS, FULLCAP, VIS>;
    smem += (size_t)WARPS_PER_CTA * VIS * sizeof(unsigned short);
    DRM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0, sms = 0;
    DRM_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WARPS_PER_CTA * 32, smem));
    DRM_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    if (per_sm < 1) return fail(DRM_EUNSUPPORTED, "search kernel does not fit on an SM with these parameters");
    long long want = (p.n + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
    long long grid = (long long)per_sm * sms;
    if (want < grid) grid = want;
    kern<<<(unsigned)grid, WARPS_PER_CTA * 32, smem, stream>>>(p);
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

template <int CS, int NS, bool FULLCAP>
int launch_reg_tf(const SearchParams &p, size_t smem, int device, cudaStream_t stream)
{
    // the visited cache's 16-bit tags are exact only below slots * 65535 vertices; DRM_SEARCH_VIS = 0 / 512 / 1024 picks (A/B)
    const char *e = getenv("DRM_SEARCH_VIS");
    int slots = e ? atoi(e) : 512;  // 512: 6 CTAs/SM, 7.5 % faster than 1024 (profiles/r1_ab_v5_search_vis_prefetch.txt)
    if (slots == 1) slots = 512;
    if (slots != 512 && slots != 1024) slots = 0;
    if (slots > 0 && p.ntotal >= (long long)slots * 65535) slots = p.ntotal < 1024ll * 65535 ? 1024 : 0;
    if (slots == 1024) return launch_reg_tfv<CS, NS, FULLCAP, 1024>(p, smem, device, stream);
    if (slots == 512) return launch_reg_tfv<CS, NS, FULLCAP, 512>(p, smem, device, stream);
    return launch_reg_tfv<CS, NS, FULLCAP, 0>(p, smem, device, stream);
}

template <int CS, int NS>
int launch_reg_t(const SearchParams &p, size_t smem, int device, cudaStream_t stream)
{
    return p.cap == 32 * NS ? launch_reg_tf<CS, NS, true>(p, smem, device, stream) : launch_reg_tf<CS, NS, false>(p, smem, device, stream);
}

template <int CS, int NPL>
int launch_t(const SearchParams &p, size_t smem, int device, cudaStream_t stream)
{
    auto kern = hnswpq_search_kernel<CS, NPL>;
    DRM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0, sms = 0;
    DRM_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WARPS_PER_CTA * 32, smem));
    DRM_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    if (per_sm < 1) return fail(DRM_EUNSUPPORTED, "search kernel does not fit on an SM with these parameters");
    long long want = (p.n + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
    long long grid = (long long)per_sm * sms;
    if (want < grid) grid = want;
    kern<<<(unsigned)grid, WARPS_PER_CTA * 32, smem, stream>>>(p);
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

}  // namespace

// Scratch for the work counter: one unsigned per device, allocated lazily.
// Work counters: a ring of 1024 slots per device so that launches in flight on different streams never share one.
static unsigned *g_work_counter[64] = {nullptr};
static std::atomic<unsigned> g_work_slot[64];
static std::once_flag g_work_once[64];

int launch_search(const DeviceIndex &dx, const float *q_dev, int64_t n, int k, int ef, float *D_dev, int64_t *I_dev,
                  int64_t *counters_dev, cudaStream_t stream)
{
    if (n <= 0) return fail(DRM_EEMPTY, "Query data is empty");
    if (k <= 0 || ef <= 0) return fail(DRM_EINVAL, "k and ef must be positive");
    if (n > 0xfffffff0ll) return fail(DRM_EINVAL, "more than 2^32 queries in one call");
    // Default: the FAISS-exact kernel (search_exact.cu: FAISS's own tie behaviour).  DRM_SEARCH_KERNEL = reg | smem | pair selects
    // one of the canonical-order kernels below (ties ordered by (distance, id); equal to FAISS up to distance ties) for A/B runs;
    // shapes outside the exact kernel (k > 256) also land here.
    {
        const char *force = getenv("DRM_SEARCH_KERNEL");
        const bool canonical = force && (!strcmp(force, "reg") || !strcmp(force, "smem") || !strcmp(force, "pair"));
        if (!canonical && search_exact_supports(dx, k, ef)) return launch_search_exact(dx, q_dev, n, k, ef, D_dev, I_dev, counters_dev, stream);
    }
    if (dx.ksub != 256) return fail(DRM_EUNSUPPORTED, "only nbits == 8 product quantizers are supported by the search kernel");
    if (dx.W0 > 64 || dx.Wu > 32) return fail(DRM_EUNSUPPORTED, "M_hnsw > 32 is not supported by the search kernel");
    if (dx.code_size != 8 && dx.code_size != 16) return fail(DRM_EUNSUPPORTED, "only M_pq of 8 or 16 (nbits 8) is supported by the search kernel");
    const int cap = ef > k ? ef : k;              // logical capacity: exactly FAISS's max(efSearch, k)
    const int cap_alloc = (cap + 31) / 32 * 32;
    if (cap > 2048) return fail(DRM_EUNSUPPORTED, "max(ef, k) above 2048 is not supported");
    if (dx.device < 0 || dx.device >= 64) return fail(DRM_EINVAL, "bad device ordinal");
    {
        cudaError_t aerr = cudaSuccess;
        std::call_once(g_work_once[dx.device], [&] { aerr = cudaMalloc(&g_work_counter[dx.device], 1024 * sizeof(unsigned)); });
        if (aerr != cudaSuccess || !g_work_counter[dx.device]) return fail(DRM_ECUDA, "cudaMalloc of the search work counters failed");
    }
    unsigned *work = g_work_counter[dx.device] + (g_work_slot[dx.device].fetch_add(1) & 1023u);
    DRM_CUDA_TRY(cudaMemsetAsync(work, 0, sizeof(unsigned), stream));

    SearchParams p{};
    p.q = q_dev; p.n = n; p.d = dx.d; p.M = dx.M; p.dsub = dx.dsub;
    p.centroids = dx.centroids; p.centroids_t = reinterpret_cast<const float4 *>(dx.centroids_t); p.rows0 = dx.rows0; p.up = dx.up;
    p.row0_bytes = dx.row0_bytes; p.rowu_bytes = dx.rowu_bytes; p.W0 = dx.W0; p.Wu = dx.Wu; p.c0_off = dx.code0_off; p.cu_off = dx.codeu_off;
    p.entry_vertex = dx.entry_vertex; p.entry_row = dx.entry_row; p.ntotal = dx.ntotal; p.max_level = dx.max_level;
    p.k = k; p.ef = ef; p.cap = cap; p.cap_alloc = cap_alloc;
    p.prefetch_depth = 0;
    if (const char *e = getenv("DRM_SEARCH_PF")) p.prefetch_depth = atoi(e) < 0 ? 0 : atoi(e) > 2 ? 2 : atoi(e);
    p.D = D_dev; p.I = reinterpret_cast<long long *>(I_dev);
    p.counters = reinterpret_cast<u64 *>(counters_dev);
    p.next_query = work;
    for (int i = 0; i < 64; ++i) p.entry_code[i] = dx.entry_code[i];
    // Kernel choice: the register-pool kernel when its preconditions hold (DRM_SEARCH_KERNEL=smem forces the other one,
    // for A/B measurements and the parity tests); both implement the same rule and return identical bits.
    const char *force = getenv("DRM_SEARCH_KERNEL");
    const bool want_smem = force && !strcmp(force, "smem");
    if (!want_smem && dx.W0 <= 32 && k <= ef && cap <= 256) {
        const size_t smem_r = (size_t)WARPS_PER_CTA * (size_t)dx.M * 256 * sizeof(float);
        const int ns = cap <= 64 ? 2 : cap <= 128 ? 4 : 8;
        if (force && !strcmp(force, "pair")) {  // two rows per iteration (A/B; same bits)
            if (dx.code_size == 8) return ns == 2 ? launch_pair_t<8, 2>(p, smem_r, dx.device, stream) : ns == 4 ? launch_pair_t<8, 4>(p, smem_r, dx.device, stream) : launch_pair_t<8, 8>(p, smem_r, dx.device, stream);
            return ns == 2 ? launch_pair_t<16, 2>(p, smem_r, dx.device, stream) : ns == 4 ? launch_pair_t<16, 4>(p, smem_r, dx.device, stream) : launch_pair_t<16, 8>(p, smem_r, dx.device, stream);
        }
        if (dx.code_size == 8) return ns == 2 ? launch_reg_t<8, 2>(p, smem_r, dx.device, stream) : ns == 4 ? launch_reg_t<8, 4>(p, smem_r, dx.device, stream) : launch_reg_t<8, 8>(p, smem_r, dx.device, stream);
        return ns == 2 ? launch_reg_t<16, 2>(p, smem_r, dx.device, stream) : ns == 4 ? launch_reg_t<16, 4>(p, smem_r, dx.device, stream) : launch_reg_t<16, 8>(p, smem_r, dx.device, stream);
    }
    size_t smem = (size_t)WARPS_PER_CTA * ((size_t)dx.M * 256 * sizeof(float) + (size_t)cap_alloc * sizeof(u64));
    const bool wide = dx.W0 > 32;
    if (dx.code_size == 8) return wide ? launch_t<8, 2>(p, smem, dx.device, stream) : launch_t<8, 1>(p, smem, dx.device, stream);
    return wide ? launch_t<16, 2>(p, smem, dx.device, stream) : launch_t<16, 1>(p, smem, dx.device, stream);
}

}  // namespace drm
