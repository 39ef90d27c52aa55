// HNSW-PQ search on sm_100a with FAISS's OWN tie behaviour: one warp per query.
//
// Replaces faiss::IndexHNSWPQ::search as called by the reference's faiss_search (src/hnswpq/search.cpp:13,39-40) and
// follows oracle/hnswpq_oracle.cpp search_one(tie_rule = 0) - the restatement of FAISS HNSW::search_from_candidates -
// decision for decision, INCLUDING what happens when two PQ distances are bit-equal (frequent: overlapping windows share
// PQ codes).  The canonical-order kernels of search_kernel.cu agree with FAISS only up to such ties, and a tie resolved
// differently changes which vertex is expanded next, so 1-3 % of result rows end up differing by more than a tie.  Here:
//
//   * the candidate MinimaxHeap is kept in shared memory in FAISS's binary max-heap ARRAY LAYOUT (1-based, entry j =
//     {distance bits, id + 1 | popped << 31}), because pop_min breaks distance ties by the highest array slot and the slot of an entry is
//     the product of every heap_push / heap_pop before it (faiss/utils/Heap.h, CMax<float, int> with the id tie-break cmp2;
//     a popped entry compares with id -1 like in FAISS, but keeps its id for the duplicate test);
//       - push (heap not full): the sift-up path is the new leaf's ancestor chain, known in advance: lane j loads ancestor j,
//         one ballot gives the number of ancestors that move down, one predicated store each;
//       - push (heap full): `if (d >= dis[0]) return`, heap_pop (the last entry sifts down from the root; warp-uniform loop,
//         one LDS.128 fetches both children) then the sift-up above;
//       - pop_min: per-lane scan + REDUX.MIN on the distance bits + REDUX.MAX on the slot;
//       - count_below(d0) >= efSearch: only reachable when k > efSearch (capacity = max(efSearch, k));
//   * the result heap has layout-free semantics (accept iff d < current worst distance, evict the largest (d, id), output
//     ascending (d, id) with (FLT_MAX, -1) padding), so it is a sorted list in REGISTERS, NSR entries per lane, split into
//     a distance word and an id word (32-bit compares, no 64-bit selects);
//   * FAISS's VisitedTable: a vertex evaluated before is either still in the candidate array (live or popped - the
//     duplicate test) or was rejected / evicted at a distance >= the current root, which only decreases, so `d >= dis[0]`
//     rejects it again, and the result heap (worst distance <= the candidate root once both are full) rejects it too.  The
//     per-warp 16-bit tag cache of search_kernel.cu is kept as a filter in front (a hit = evaluated before).
//
// Memory: one hop = one contiguous fused row (ids + the PQ codes of those neighbours, index_image.cu); ADC table in
// shared memory; fp32 arithmetic in the oracle's order with round-to-nearest intrinsics (bit-identical distances).
#include "common.h"

#include <atomic>
#include <cfloat>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace drm {

namespace {

constexpr int XWARPS = 4;
constexpr unsigned FULL = 0xffffffffu;
constexpr unsigned POPPED = 0x80000000u;
constexpr unsigned R_EMPTY_D = 0x7f7fffffu;  // FLT_MAX: FAISS initialises the result heap to (FLT_MAX, -1)
constexpr unsigned R_EMPTY_I = 0xffffffffu;
typedef unsigned long long u64;

struct ExactParams {
    const float *q;
    long long n;
    int d, M, dsub;
    const float *centroids;
    const float4 *centroids_t;
    const uint8_t *rows0;
    const uint8_t *up;
    int row0_bytes, rowu_bytes, W0, Wu, c0_off, cu_off;
    long long entry_vertex, entry_row, ntotal;
    int max_level;
    int k, ef, cap;
    int c_words;          // 32-bit words of shared memory per warp for the candidate heap
    int readahead;        // 1: read the row of the runner-up candidate into registers while this row is processed
    int prefetch_accept;  // 1: L2-prefetch the row of every vertex accepted into the candidate heap; 1 + s: only of those in the lowest
                          //    2^-s of the [popped distance, heap root] interval
    int prefetch_l1;      // 1: the accept-time prefetch targets L1 instead of L2 (DRM_SEARCH_PFL1, A/B)
    int tma;              // 1: rows arrive through the bulk-copy engine (cp.async.bulk + mbarrier) into a per-warp staging buffer,
                          //    and the accept-time prefetch is one cp.async.bulk.prefetch.L2 instead of one prefetch per line
    int row_stage_bytes;  // per-warp staging: row bytes rounded to 16, + 16 for the mbarrier (0 when tma is off)
    float *D;
    long long *I;
    u64 *counters;
    unsigned *next_query;
    long long id_add;     // added to every filled label (shard-local -> global window id)
    uint8_t entry_code[64];
    SearchDispatch disp;  // world == 0: results go to D / I rows of the launch
};

template <int CS> struct XCode;
template <> struct XCode<8> { uint2 v; };
template <> struct XCode<16> { uint4 v; };
template <int CS> __device__ __forceinline__ XCode<CS> x_load_code(const uint8_t *p);
template <> __device__ __forceinline__ XCode<8> x_load_code<8>(const uint8_t *p)
{
    XCode<8> c; c.v = __ldg(reinterpret_cast<const uint2 *>(p)); return c;
}
template <> __device__ __forceinline__ XCode<16> x_load_code<16>(const uint8_t *p)
{
    XCode<16> c; c.v = __ldg(reinterpret_cast<const uint4 *>(p)); return c;
}
__device__ __forceinline__ float x_adc_word(const float *T, unsigned w, int m0, float acc, bool first)
{
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        float t = T[(m0 + b) * 256 + ((w >> (8 * b)) & 0xffu)];
        acc = (first && b == 0) ? t : __fadd_rn(acc, t);
    }
    return acc;
}
__device__ __forceinline__ float x_adc(const float *T, const XCode<8> &c)
{
    float a = x_adc_word(T, c.v.x, 0, 0.f, true);
    return x_adc_word(T, c.v.y, 4, a, false);
}
__device__ __forceinline__ float x_adc(const float *T, const XCode<16> &c)
{
    float a = x_adc_word(T, c.v.x, 0, 0.f, true);
    a = x_adc_word(T, c.v.y, 4, a, false);
    a = x_adc_word(T, c.v.z, 8, a, false);
    return x_adc_word(T, c.v.w, 12, a, false);
}
template <int CS> __device__ __forceinline__ float x_adc_bytes(const float *T, const uint8_t *code)
{
    float a = T[code[0]];
    for (int m = 1; m < CS; ++m) a = __fadd_rn(a, T[m * 256 + code[m]]);
    return a;
}

// T[m][j] = sum_t (x[m*dsub+t] - c[m][j][t])^2, t ascending, one fma per term (oracle adc_table)
__device__ __forceinline__ void x_build_lut(float *T, const float *__restrict__ x, const float *__restrict__ cent, const float4 *__restrict__ cent_t,
                                            int M, int dsub, int lane)
{
    if (cent_t) {
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
        return;
    }
    for (int m = 0; m < M; ++m) {
        const float *xm = x + m * dsub;
        for (int j = lane; j < 256; j += 32) {
            const float *c = cent + ((size_t)(m * 256 + j)) * dsub;
            float acc = 0.f;
            for (int t = 0; t < dsub; ++t) {
                float e = __fsub_rn(__ldg(xm + t), __ldg(c + t));
                acc = __fmaf_rn(e, e, acc);
            }
            T[m * 256 + j] = acc;
        }
    }
}

__device__ __forceinline__ u64 x_warp_min_u64(u64 v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        u64 other = __shfl_xor_sync(FULL, v, o);
        v = other < v ? other : v;
    }
    return v;
}

// FAISS cmp2 order of a candidate-heap entry: (distance, id) with a popped entry's id read as -1.  The stored word is
// id + 1 (>= 1) with bit 31 set once popped, so max((int)w, 0) is "effective id + 1".
__device__ __forceinline__ u64 ckey(unsigned d, unsigned w)
{
    const int e = max((int)w, 0);
    return ((u64)d << 32) | (unsigned)e;
}

// MinimaxHeap::pop_min's choice: smallest distance among live entries, highest slot among equal minima.  Returns the slot
// (0 when no live entry) and the distance bits.  CL > 0: the capacity is at most 32 * CL and the scan is unrolled.
template <int CL>
__device__ __forceinline__ unsigned c_find_min(const uint2 *C2, int csize, int lane, unsigned &dmin)
{
    unsigned bd = 0xffffffffu, bj = 0;
    if (CL > 0) {
#pragma unroll
        for (int t = 0; t < (CL > 0 ? CL : 1); ++t) {
            const int j = lane + 1 + 32 * t;
            if (j <= csize) {
                const uint2 e = C2[j];
                if ((int)e.y > 0 && e.x <= bd) { bd = e.x; bj = (unsigned)j; }
            }
        }
    } else {
        for (int j = lane + 1; j <= csize; j += 32) {
            const uint2 e = C2[j];
            if ((int)e.y > 0 && e.x <= bd) { bd = e.x; bj = (unsigned)j; }
        }
    }
    dmin = __reduce_min_sync(FULL, bd);
    return __reduce_max_sync(FULL, bd == dmin ? bj : 0u);
}

// VisitedTable back-stop: is vertex word `w` (id + 1) in the candidate array, live or popped?
template <int CL>
__device__ __forceinline__ bool c_contains(const uint2 *C2, int csize, int lane, unsigned w)
{
    bool dup = false;
    if (CL > 0) {
#pragma unroll
        for (int t = 0; t < (CL > 0 ? CL : 1); ++t) {
            const int j = lane + 1 + 32 * t;
            if (j <= csize) dup = dup || (C2[j].y & ~POPPED) == w;
        }
    } else {
        for (int j = lane + 1; j <= csize; j += 32) dup = dup || (C2[j].y & ~POPPED) == w;
    }
    return __any_sync(FULL, dup);
}

// faiss::heap_push<CMax<float,int>>(k, ...) with k the size AFTER the push: the new entry climbs from slot k while it is
// greater (cmp2) than its parent.  Lane j (>= 1) owns ancestor k >> j.
__device__ __forceinline__ void c_heap_push(uint2 *C2, int k, unsigned kd, unsigned kw, int lane)
{
    const int a = k >> lane;
    const bool has = lane >= 1 && a >= 1;
    uint2 e = make_uint2(0u, 0u);
    if (has) e = C2[a];
    const bool gt = has && ckey(kd, kw) > ckey(e.x, e.y);
    const unsigned m = __ballot_sync(FULL, gt) >> 1;  // bit j-1: the new entry passes ancestor j
    const int cnt = __ffs((int)~m) - 1;                // FAISS stops at the first ancestor it does not pass
    __syncwarp();
    if (lane >= 1 && lane <= cnt) C2[k >> (lane - 1)] = e;
    if (lane == 0) C2[k >> cnt] = make_uint2(kd, kw);
    __syncwarp();
}

// faiss::heap_pop<CMax<float,int>>(k, ...): the root leaves, entry k sifts down from the root along the path of larger
// children until it is greater than the larger child.  That path does not depend on the sinking entry, so instead of a
// 7-deep chain of dependent shared-memory reads: (1) every lane computes, for its share of the internal nodes, which
// child is larger (one LDS.128 per node, two ballots); (2) the path is read off those bits with register arithmetic;
// (3) lane j loads path node j, one ballot finds where FAISS's loop stops, and the nodes above it move up one level with one
// predicated store each.  Same comparisons as FAISS on the same array (entry k itself stays in place while it is compared,
// as in FAISS, where `i2 == k + 1` and cmp2(v, v) = false make it harmless).  NB = ballot words = ceil(cap / 64).
// Capacities above 256: the plain dependent chain, walked by one lane.
__device__ __forceinline__ void c_heap_pop_seq(uint2 *C2, int k, int lane)
{
    if (lane == 0) {
        const uint2 v = C2[k];
        const u64 kv = ckey(v.x, v.y);
        int i = 1;
        for (;;) {
            const int i1 = i << 1;
            if (i1 > k) break;
            const uint4 c = *reinterpret_cast<const uint4 *>(C2 + i1);
            const u64 k1 = ckey(c.x, c.y), k2 = ckey(c.z, c.w);
            const bool first = (i1 == k) || k1 > k2;
            if (kv > (first ? k1 : k2)) break;
            C2[i] = first ? make_uint2(c.x, c.y) : make_uint2(c.z, c.w);
            i = first ? i1 : i1 + 1;
        }
        C2[i] = v;
    }
    __syncwarp();
}

template <int NB>
__device__ __forceinline__ void c_heap_pop(uint2 *C2, int k, int lane)
{
    unsigned Bw[NB];
#pragma unroll
    for (int t = 0; t < NB; ++t) {
        const int i = lane + 1 + 32 * t, i1 = i << 1;
        bool first = true;
        if (i1 <= k) {
            const uint4 c = *reinterpret_cast<const uint4 *>(C2 + i1);
            first = (i1 == k) || ckey(c.x, c.y) > ckey(c.z, c.w);
        }
        Bw[t] = __ballot_sync(FULL, first);
    }
    // the path of larger children, read off the bits: node i -> 2i + (first ? 0 : 1).  Unrolled to the depth of a 64 * NB heap.
    unsigned nB[NB];
#pragma unroll
    for (int t = 0; t < NB; ++t) nB[t] = ~Bw[t];
    constexpr int DEPTH = NB == 1 ? 6 : NB == 2 ? 7 : 8;
    int i = 1, mine = 1, L = 0;
#pragma unroll
    for (int j = 1; j <= DEPTH; ++j) {
        if ((i << 1) <= k) {
            unsigned w = nB[0];
            if (j > 5) {  // a node at depth j - 1 <= 4 is one of the first 31: its bit is in word 0 (j is a compile-time constant here)
#pragma unroll
                for (int t = 1; t < NB; ++t)
                    if (((i - 1) >> 5) == t) w = nB[t];
            }
            i = (i << 1) + (int)((w >> ((i - 1) & 31)) & 1u);
            if (lane == j) mine = i;
            L = j;
        }
    }
    const uint2 v = C2[k];
    uint2 e = make_uint2(0u, 0u);
    if (lane >= 1 && lane <= L) e = C2[mine];
    const bool gt = lane >= 1 && lane <= L && ckey(v.x, v.y) > ckey(e.x, e.y);
    const unsigned m = __ballot_sync(FULL, gt);
    const int stop = m ? __ffs((int)m) - 1 : L + 1;  // first path node (below the root) that the sinking entry beats
    __syncwarp();
    if (lane >= 1 && lane < stop) C2[mine >> 1] = e;
    if (lane == stop - 1) C2[mine] = v;
    __syncwarp();
}

// Result list: position i lives in lane i / NSR, slot i % NSR; ascending (distance, id); empty slots are (FLT_MAX, -1).
template <int NSR, bool FULLK>
__device__ __forceinline__ void r_insert(unsigned (&rh)[NSR], unsigned (&rl)[NSR], unsigned &rthr, unsigned kd, unsigned kid, int lane, int k,
                                         int max_lane, int max_slot)
{
    const u64 kin = ((u64)kd << 32) | kid;
    bool lt[NSR];
#pragma unroll
    for (int s = 0; s < NSR; ++s) lt[s] = (((u64)rh[s] << 32) | rl[s]) < kin;
    const unsigned ph = __shfl_up_sync(FULL, rh[NSR - 1], 1), pl = __shfl_up_sync(FULL, rl[NSR - 1], 1);
    const bool plt = lane == 0 || (((u64)ph << 32) | pl) < kin;
#pragma unroll
    for (int s = NSR - 1; s >= 1; --s) {
        rh[s] = lt[s] ? rh[s] : (lt[s - 1] ? kd : rh[s - 1]);
        rl[s] = lt[s] ? rl[s] : (lt[s - 1] ? kid : rl[s - 1]);
    }
    rh[0] = lt[0] ? rh[0] : (plt ? kd : ph);
    rl[0] = lt[0] ? rl[0] : (plt ? kid : pl);
    if (FULLK) {
        rthr = __shfl_sync(FULL, rh[NSR - 1], 31);
    } else {
        const int base = lane * NSR;
        unsigned last = rh[0];
#pragma unroll
        for (int s = 0; s < NSR; ++s) {
            if (base + s >= k) { rh[s] = R_EMPTY_D; rl[s] = R_EMPTY_I; }
            if (s == max_slot) last = rh[s];
        }
        rthr = __shfl_sync(FULL, last, max_lane);
    }
}

// CL: the candidate capacity is at most 32 * CL (scans unrolled, warp-parallel sift-down); 0 = any capacity up to 2048.
template <int CS, int NSR, bool FULLK, int VIS, int CL>
__global__ void __launch_bounds__(XWARPS * 32, VIS ? 5 : 6) hnswpq_search_exact_kernel(const ExactParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int VIS_SLOTS = VIS > 0 ? VIS : 1;
    const size_t t_bytes = (size_t)p.M * 256 * sizeof(float);
    constexpr int BLOOM_WORDS = VIS > 0 ? 32 : 0;  // 1024-bit filter of the vertices the tag cache has lost
    const size_t per_warp = t_bytes + (size_t)p.c_words * 4 + (size_t)VIS * 2 + BLOOM_WORDS * 4 + (size_t)p.row_stage_bytes;
    float *T = reinterpret_cast<float *>(smem_raw + warp * per_warp);
    uint2 *C2 = reinterpret_cast<uint2 *>(smem_raw + warp * per_warp + t_bytes);  // entry j at C2[j], j = 1 .. cap (+1 spare)
    unsigned short *vis = reinterpret_cast<unsigned short *>(smem_raw + warp * per_warp + t_bytes + (size_t)p.c_words * 4);
    unsigned *bloom = reinterpret_cast<unsigned *>(smem_raw + warp * per_warp + t_bytes + (size_t)p.c_words * 4 + (size_t)VIS * 2);
    // bulk-copy row staging (p.tma): [row bytes][8-byte mbarrier]; one copy in flight per warp, phase bit flips per use
    unsigned char *rowbuf = smem_raw + warp * per_warp + t_bytes + (size_t)p.c_words * 4 + (size_t)VIS * 2 + BLOOM_WORDS * 4;
    unsigned tma_phase = 0;
#ifndef DRM_EMU
    const unsigned rowbuf_s = (unsigned)__cvta_generic_to_shared(rowbuf);
    const unsigned bar_s = rowbuf_s + (unsigned)p.row_stage_bytes - 16u;
    if (p.tma) {
        if (lane == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
    }
#endif
    const int cap = p.cap, k = p.k;
    const int max_lane = (k - 1) / NSR, max_slot = (k - 1) % NSR;

    for (;;) {
        unsigned qi = 0;
        if (lane == 0) qi = atomicAdd(p.next_query, 1u);
        qi = __shfl_sync(FULL, qi, 0);
        if ((long long)qi >= p.n) break;

        __syncwarp();
        x_build_lut(T, p.q + (size_t)qi * p.d, p.centroids, p.centroids_t, p.M, p.dsub, lane);
        if (VIS) {
            uint4 *v4 = reinterpret_cast<uint4 *>(vis);
#pragma unroll
            for (int i = 0; i < (VIS_SLOTS * 2 / 16 + 31) / 32; ++i)
                if (lane + 32 * i < VIS_SLOTS * 2 / 16) v4[lane + 32 * i] = make_uint4(~0u, ~0u, ~0u, ~0u);
            bloom[lane] = 0u;
        }
        __syncwarp();

        unsigned nhops0 = 0, nhops_up = 0, ndis = 0;

        // ---------------- greedy descent, levels max_level .. 1 (FAISS greedy_update_nearest)
        float d_near = x_adc_bytes<CS>(T, p.entry_code);
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
                    if (lane < first_neg) dj = x_adc(T, x_load_code<CS>(row + p.cu_off + CS * lane));
                    u64 best = x_warp_min_u64(((u64)__float_as_uint(dj) << 32) | (unsigned)lane);
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

        // ---------------- level 0 (HNSW::search_from_candidates)
        unsigned rh[NSR], rl[NSR];
#pragma unroll
        for (int s = 0; s < NSR; ++s) { rh[s] = R_EMPTY_D; rl[s] = R_EMPTY_I; }
        unsigned rthr = R_EMPTY_D;
        int csize = 0, nvalid = 0;
        {
            const unsigned kd = __float_as_uint(d_near), kid = (unsigned)vertex;
            c_heap_push(C2, ++csize, kd, kid + 1u, lane);
            ++nvalid;
            if (kd < rthr) r_insert<NSR, FULLK>(rh, rl, rthr, kd, kid, lane, k, max_lane, max_slot);
            if (VIS && lane == 0) vis[kid % VIS_SLOTS] = (unsigned short)(kid / VIS_SLOTS);
        }
        unsigned spec_v = 0xffffffffu;
        int nid_s = -1;
        XCode<CS> code_s;
        code_s.v = {};
        while (nvalid > 0) {
            unsigned d0;
            const unsigned jm = c_find_min<CL>(C2, csize, lane, d0);
            const unsigned w0 = C2[jm].y;
            const unsigned v0 = (w0 & ~POPPED) - 1u;
            __syncwarp();
            if (lane == 0) C2[jm].y = w0 | POPPED;
            __syncwarp();
            --nvalid;
            if (cap > p.ef) {  // MinimaxHeap::count_below(d0) >= efSearch: popped entries keep their distance and count
                unsigned below = 0;
                for (int j = lane + 1; j <= csize; j += 32) below += C2[j].x < d0;
                if ((int)__reduce_add_sync(FULL, below) >= p.ef) break;
            }
            int nid; XCode<CS> code;
            if (v0 == spec_v) { nid = nid_s; code = code_s; }
#ifndef DRM_EMU
            else if (p.tma) {
                // the whole fused row (contiguous, 16-byte aligned: the regular access of this kernel) in ONE bulk copy issued by one lane
                const uint8_t *row = p.rows0 + (size_t)v0 * p.row0_bytes;
                if (lane == 0) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the buffer was last touched by ordinary loads
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(p.row0_bytes) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(rowbuf_s), "l"(row),
                                 "r"(p.row0_bytes), "r"(bar_s)
                                 : "memory");
                }
                unsigned done = 0;
                do {
                    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                                 : "=r"(done)
                                 : "r"(bar_s), "r"(tma_phase)
                                 : "memory");
                } while (!done);
                tma_phase ^= 1u;
                nid = lane < p.W0 ? reinterpret_cast<const int *>(rowbuf)[lane] : -1;
                code.v = {};
                if (lane < p.W0) {
                    if (CS == 8) { const uint2 t = *reinterpret_cast<const uint2 *>(rowbuf + p.c0_off + 8 * lane); code.v.x = t.x; code.v.y = t.y; }
                    else code = *reinterpret_cast<const XCode<CS> *>(rowbuf + p.c0_off + CS * lane);
                }
                __syncwarp();  // every lane has read its slot before the next copy may land
            }
#endif
            else {
                const uint8_t *row = p.rows0 + (size_t)v0 * p.row0_bytes;
                nid = lane < p.W0 ? __ldg(reinterpret_cast<const int *>(row) + lane) : -1;
                code.v = {};
                if (lane < p.W0) code = x_load_code<CS>(row + p.c0_off + CS * lane);
            }
            if (p.readahead) {  // the row pop_min would take next if this row inserted nothing
                unsigned d1;
                const unsigned j1 = c_find_min<CL>(C2, csize, lane, d1);
                if (j1) {
                    spec_v = (C2[j1].y & ~POPPED) - 1u;
                    const uint8_t *nrow = p.rows0 + (size_t)spec_v * p.row0_bytes;
                    nid_s = lane < p.W0 ? __ldg(reinterpret_cast<const int *>(nrow) + lane) : -1;
                    if (lane < p.W0) code_s = x_load_code<CS>(nrow + p.c0_off + CS * lane);
                } else spec_v = 0xffffffffu;
            }
            ++nhops0;

            const unsigned neg = __ballot_sync(FULL, nid < 0);
            const int first_neg = neg ? (__ffs(neg) - 1) : 32;
            bool valid = lane < first_neg;
            // at_risk: this vertex may have been evaluated before although the tag cache does not say so (its tag was
            // overwritten by a colliding vertex, or it lost a same-hop race for the slot) - only those need the array scan
            bool at_risk = VIS == 0;
            if (VIS) {
                const unsigned slot = (unsigned)nid % VIS_SLOTS;
                const unsigned short tag = (unsigned short)((unsigned)nid / VIS_SLOTS);
                unsigned short old = 0xffffu;
                if (valid) {
                    old = vis[slot];
                    if (old == tag) valid = false;
                }
                const unsigned fresh = __ballot_sync(FULL, valid);  // (also: every lane has read before any lane writes)
                ndis += (unsigned)__popc(fresh);
                if (!fresh) continue;
                if (valid) vis[slot] = tag;
                __syncwarp();
                if (valid) {
                    if (old != 0xffffu) {  // a cached vertex loses its tag: remember that it was evaluated
                        const unsigned lost = (unsigned)old * VIS_SLOTS + slot;
                        const unsigned h = (lost * 2654435761u) >> 22;
                        atomicOr(&bloom[h >> 5], 1u << (h & 31));
                    }
                    if (vis[slot] != tag) {  // lost a same-hop race for the slot: this vertex itself is now uncached
                        const unsigned h = ((unsigned)nid * 2654435761u) >> 22;
                        atomicOr(&bloom[h >> 5], 1u << (h & 31));
                    }
                }
                __syncwarp();
                if (valid) {
                    const unsigned h = ((unsigned)nid * 2654435761u) >> 22;
                    at_risk = (bloom[h >> 5] >> (h & 31)) & 1u;
                }
            } else {
                ndis += (unsigned)first_neg;
            }
            unsigned dbits = 0xffffffffu;
            if (valid) dbits = __float_as_uint(x_adc(T, code));
            unsigned root_d = csize == cap ? C2[1].x : 0xffffffffu;
            // accept-time prefetch gate (prefetch_accept = 1 + s): only vertices in the lowest 2^-s of the [popped distance, root]
            // interval (on the distance bits) - those are popped soon; the rest are mostly evicted before their turn, and a row
            // prefetched for them is DRAM traffic for nothing (s = 0: every accepted vertex)
            const unsigned pf_thr = p.prefetch_accept <= 1 ? 0xffffffffu : d0 + ((root_d - d0) >> (p.prefetch_accept - 1));
            unsigned mask = __ballot_sync(FULL, valid && dbits < root_d);  // the root only decreases: a safe pre-filter
            // a vertex id that occurs twice in this row is new the first time only: the later copies take the scan
            const unsigned twin = __match_any_sync(FULL, valid ? nid : -1 - lane);
            const unsigned risk = __ballot_sync(FULL, at_risk || (twin & ((1u << lane) - 1u)) != 0u);
            while (mask) {  // neighbours in row order, each seeing the heaps as the previous one left them
                const int src = __ffs(mask) - 1;
                mask &= mask - 1;
                const unsigned kd = __shfl_sync(FULL, dbits, src);
                const unsigned kid = (unsigned)__shfl_sync(FULL, nid, src);
                const bool full = csize == cap;
                if (full && kd >= root_d) continue;  // MinimaxHeap::push: `if (v >= dis[0]) return`
                // VisitedTable: evaluated before and still in the array (live or popped)?  Also catches an id listed twice in a row.
                if ((risk >> src) & 1u) { if (c_contains<CL>(C2, csize, lane, kid + 1u)) continue; }
                if (kd < rthr) r_insert<NSR, FULLK>(rh, rl, rthr, kd, kid, lane, k, max_lane, max_slot);  // `if (d < threshold) add_result`
                if (full) {
                    const unsigned rootw = C2[1].y;
                    __syncwarp();                    // every lane has read the root before lane 0 rewrites it
                    if ((int)rootw > 0) --nvalid;    // the evicted root was still live
                    if (CL > 0) c_heap_pop<(CL > 1 ? CL / 2 : 1)>(C2, csize--, lane);
                    else c_heap_pop_seq(C2, csize--, lane);
                }
                c_heap_push(C2, ++csize, kd, kid + 1u, lane);
                ++nvalid;
                if (csize == cap) root_d = C2[1].x;
#ifndef DRM_EMU
                if (p.tma) {
                    if (p.prefetch_accept && lane == 0 && kd <= pf_thr) {
                        const uint8_t *prow = p.rows0 + (size_t)kid * p.row0_bytes;
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(prow), "r"(p.row0_bytes) : "memory");
                    }
                } else if (p.prefetch_accept && lane < 3 && kd <= pf_thr) {
                    const uint8_t *prow = p.rows0 + (size_t)kid * p.row0_bytes + lane * 128;
                    if (lane * 128 < p.row0_bytes) {
                        if (p.prefetch_l1) asm volatile("prefetch.global.L1 [%0];" ::"l"(prow));
                        else asm volatile("prefetch.global.L2 [%0];" ::"l"(prow));
                    }
                }
#endif
            }
        }

        // ---------------- results: heap_reorder = ascending (distance, id), (FLT_MAX, -1) padding
        float *Dq = p.D + (size_t)qi * p.k;
        long long *Iq = p.I + (size_t)qi * p.k;
        if (p.disp.world > 0) {  // fused exchange: the list goes to the rank that merges this read
            int o = 0;
            while (o + 1 < p.disp.world && (long long)qi >= p.disp.slice_lo[o + 1]) ++o;
            const size_t row = (size_t)p.disp.my_shard * p.disp.slice_cap + ((long long)qi - p.disp.slice_lo[o]);
            Dq = p.disp.D_peer[o] + row * p.k;
            Iq = p.disp.I_peer[o] + row * p.k;
        }
#pragma unroll
        for (int s = 0; s < NSR; ++s) {
            const int i = lane * NSR + s;
            if (i < k) {
                Dq[i] = __uint_as_float(rh[s]);
                Iq[i] = rl[s] == R_EMPTY_I ? -1ll : (long long)rl[s] + p.id_add;
            }
        }
        if (p.disp.world > 0) __threadfence_system();
        if (p.counters && lane == 0) {
            atomicAdd(p.counters + 0, (u64)nhops0);
            atomicAdd(p.counters + 1, (u64)nhops_up);
            atomicAdd(p.counters + 2, (u64)ndis);
        }
    }
}

template <int CS, int NSR, bool FULLK, int VIS, int CL>
int launch_exact_tfvc(const ExactParams &p, int device, cudaStream_t stream)
{
    auto kern = hnswpq_search_exact_kernel<CS, NSR, FULLK, VIS, CL>;
    const size_t smem = (size_t)XWARPS * ((size_t)p.M * 256 * sizeof(float) + (size_t)p.c_words * 4 + (size_t)VIS * 2 + (VIS > 0 ? 128 : 0) + (size_t)p.row_stage_bytes);
    DRM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0, sms = 0;
    DRM_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, XWARPS * 32, smem));
    DRM_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    if (per_sm < 1) return fail(DRM_EUNSUPPORTED, "search kernel does not fit on an SM with these parameters");
    // DRM_SEARCH_CTAS_PER_SM: fewer persistent CTAs per SM than fit, so that another kernel's CTAs (the SW rerank of the previous
    // chunk, on another stream) can share the SMs with this one (tools/overlap_probe.py)
    if (const char *ev = getenv("DRM_SEARCH_CTAS_PER_SM")) {
        const int cap_sm = atoi(ev);
        if (cap_sm >= 1 && cap_sm < per_sm) per_sm = cap_sm;
    }
    long long want = (p.n + XWARPS - 1) / XWARPS;
    long long grid = (long long)per_sm * sms;
    if (want < grid) grid = want;
    kern<<<(unsigned)grid, XWARPS * 32, smem, stream>>>(p);
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

template <int CS, int NSR, bool FULLK, int VIS>
int launch_exact_tfv(const ExactParams &p, int device, cudaStream_t stream)
{
    if (p.cap <= 64) return launch_exact_tfvc<CS, NSR, FULLK, VIS, 2>(p, device, stream);
    if (p.cap <= 128) return launch_exact_tfvc<CS, NSR, FULLK, VIS, 4>(p, device, stream);
    if (p.cap <= 256) return launch_exact_tfvc<CS, NSR, FULLK, VIS, 8>(p, device, stream);
    return launch_exact_tfvc<CS, NSR, FULLK, VIS, 0>(p, device, stream);
}

template <int CS, int NSR, bool FULLK>
int launch_exact_tf(const ExactParams &p, int device, cudaStream_t stream)
{
    // the visited cache's 16-bit tags are exact only below slots * 65535 vertices; DRM_SEARCH_VIS = 0 / 512 / 1024 picks (A/B)
    const char *e = getenv("DRM_SEARCH_VIS");
    int slots = e ? atoi(e) : 512;
    if (slots == 1) slots = 512;
    if (slots != 512 && slots != 1024) slots = 0;
    if (slots > 0 && p.ntotal >= (long long)slots * 65535) slots = p.ntotal < 1024ll * 65535 ? 1024 : 0;
    if (slots == 1024) return launch_exact_tfv<CS, NSR, FULLK, 1024>(p, device, stream);
    if (slots == 512) return launch_exact_tfv<CS, NSR, FULLK, 512>(p, device, stream);
    return launch_exact_tfv<CS, NSR, FULLK, 0>(p, device, stream);
}

template <int CS>
int launch_exact_t(const ExactParams &p, int device, cudaStream_t stream)
{
    const int k = p.k;
    if (k <= 64) return k == 64 ? launch_exact_tf<CS, 2, true>(p, device, stream) : launch_exact_tf<CS, 2, false>(p, device, stream);
    if (k <= 128) return k == 128 ? launch_exact_tf<CS, 4, true>(p, device, stream) : launch_exact_tf<CS, 4, false>(p, device, stream);
    return k == 256 ? launch_exact_tf<CS, 8, true>(p, device, stream) : launch_exact_tf<CS, 8, false>(p, device, stream);
}

}  // namespace

static unsigned *g_xwork[64] = {nullptr};
static std::atomic<unsigned> g_xslot[64];
static std::once_flag g_xonce[64];

bool search_exact_supports(const DeviceIndex &dx, int k, int ef)
{
    const int cap = ef > k ? ef : k;
    return dx.ksub == 256 && dx.W0 <= 32 && dx.Wu <= 32 && (dx.code_size == 8 || dx.code_size == 16) && k <= 256 && cap <= 2048;
}

int launch_search_exact(const DeviceIndex &dx, const float *q_dev, int64_t n, int k, int ef, float *D_dev, int64_t *I_dev,
                        int64_t *counters_dev, cudaStream_t stream)
{
    return launch_search_exact_ex(dx, q_dev, n, k, ef, D_dev, I_dev, counters_dev, stream, 0, nullptr);
}

int launch_search_exact_ex(const DeviceIndex &dx, const float *q_dev, int64_t n, int k, int ef, float *D_dev, int64_t *I_dev,
                           int64_t *counters_dev, cudaStream_t stream, int64_t id_add, const SearchDispatch *disp)
{
    if (n <= 0) return fail(DRM_EEMPTY, "Query data is empty");
    if (k <= 0 || ef <= 0) return fail(DRM_EINVAL, "k and ef must be positive");
    if (n > 0xfffffff0ll) return fail(DRM_EINVAL, "more than 2^32 queries in one call");
    if (!search_exact_supports(dx, k, ef)) return fail(DRM_EUNSUPPORTED, "search shape outside the FAISS-exact kernel (nbits 8, M_pq 8/16, M_hnsw <= 16, k <= 256, max(ef, k) <= 2048)");
    if (dx.device < 0 || dx.device >= 64) return fail(DRM_EINVAL, "bad device ordinal");
    cudaError_t aerr = cudaSuccess;
    std::call_once(g_xonce[dx.device], [&] { aerr = cudaMalloc(&g_xwork[dx.device], 1024 * sizeof(unsigned)); });
    if (aerr != cudaSuccess || !g_xwork[dx.device]) return fail(DRM_ECUDA, "cudaMalloc of the search work counters failed");
    unsigned *work = g_xwork[dx.device] + (g_xslot[dx.device].fetch_add(1) & 1023u);
    DRM_CUDA_TRY(cudaMemsetAsync(work, 0, sizeof(unsigned), stream));

    ExactParams p{};
    p.q = q_dev; p.n = n; p.d = dx.d; p.M = dx.M; p.dsub = dx.dsub;
    p.centroids = dx.centroids; p.centroids_t = reinterpret_cast<const float4 *>(dx.centroids_t); p.rows0 = dx.rows0; p.up = dx.up;
    p.row0_bytes = dx.row0_bytes; p.rowu_bytes = dx.rowu_bytes; p.W0 = dx.W0; p.Wu = dx.Wu; p.c0_off = dx.code0_off; p.cu_off = dx.codeu_off;
    p.entry_vertex = dx.entry_vertex; p.entry_row = dx.entry_row; p.ntotal = dx.ntotal; p.max_level = dx.max_level;
    p.k = k; p.ef = ef; p.cap = ef > k ? ef : k;
    p.c_words = ((p.cap + 2) * 2 + 3) / 4 * 4;  // entries 0 .. cap + 1, 8 bytes each, 16-byte multiple
    p.readahead = 0; p.prefetch_accept = 2;  // A/B on B200 (profiles/r2_ab_search_exact.txt, r2_ab_search_prefetch_gate.txt): the gated accept-time
                                             // L2 prefetch beats the register read-ahead, and the ungated one costs 20 % more DRAM traffic for nothing
    if (const char *e = getenv("DRM_SEARCH_RA")) p.readahead = atoi(e) != 0;
    if (const char *e = getenv("DRM_SEARCH_PFA")) p.prefetch_accept = atoi(e) < 0 ? 0 : atoi(e) > 8 ? 8 : atoi(e);
    p.prefetch_l1 = 0;
    if (const char *e = getenv("DRM_SEARCH_PFL1")) p.prefetch_l1 = atoi(e) != 0;
    p.tma = 0;
    if (const char *e = getenv("DRM_SEARCH_TMA")) p.tma = atoi(e) != 0;
    if (p.row0_bytes % 16) p.tma = 0;  // the bulk-copy engine moves multiples of 16 bytes
    p.row_stage_bytes = p.tma ? (p.row0_bytes + 15) / 16 * 16 + 16 : 0;
    p.D = D_dev; p.I = reinterpret_cast<long long *>(I_dev);
    p.counters = reinterpret_cast<u64 *>(counters_dev);
    p.next_query = work;
    p.id_add = id_add;
    if (disp) p.disp = *disp;
    for (int i = 0; i < 64; ++i) p.entry_code[i] = dx.entry_code[i];
    return dx.code_size == 8 ? launch_exact_t<8>(p, dx.device, stream) : launch_exact_t<16>(p, dx.device, stream);
}

}  // namespace drm
