// Smith-Waterman rerank on sm_100a.
//
// Replaces post_process_sw_static / post_process_sw_dynamic (src/utils/post_processor.cpp:357-549) with
// its callees find_sequences (:74-336), find_sequence (:47-66), sw_reranker (src/utils/reranker.cpp:3-51)
// and calc_sw_score (src/utils/metrics.cpp:10-45): match +1, mismatch -1, linear gap -1, byte equality,
// score = max over all cells.
//
// Fast kernel (sw_rerank_kernel): one CTA per read.  Candidate windows are scored two at a time in
// the two 16-bit halves of a register (inter-task SIMD, native VIMNMX3.S16x2 / VIADDMNMX.S16x2.RELU on
// sm_100a), each pair by a QUAD of lanes: lane l of the quad owns query columns [38l, 38l+38) in
// registers and the window rows flow through the quad as a 4-deep software pipeline (one shuffle per
// row).  Per cell pair:  s2 = PRMT(profile[j], sel_row)   (0 / 2 for mismatch / match)
//                        m  = max3(diag + s2, up, left)   (= H + 1 before clamping)
//                        h  = max(m - 1, 0)
//                        best = max3(best, m_j, m_j+1)    (every other column)
// The genome is held as 1-byte base classes (A,C,G,T,N = 0..4); reverse-complement windows are
// produced by index arithmetic (3 - c), never materialised.
//
// Generic kernel (sw_generic_kernel): one thread per pair, int32, raw byte equality; used for
// drm_sw_pairs and whenever the fast path's preconditions fail (query longer than 152, ref_len above
// 256, bytes outside ACGTN in the reference).  Both are CUDA; there is no host path.
#include "common.h"

#include <cuda_fp16.h>

#include <cstdlib>
#include <cstring>

namespace drm {

namespace {

typedef unsigned long long u64;
constexpr unsigned FULL = 0xffffffffu;
constexpr int CPL = 38;             // query columns per lane
constexpr int QCOLS = 4 * CPL;      // 152 = 150 + the '<' '>' decoration of format_fastq
constexpr int MAX_FAST_REFLEN = 256;
constexpr int MAXC = 1024;          // candidates per read handled in shared memory
constexpr int MAXTAKE = 1024;
constexpr int RERANK_THREADS = 256;
constexpr int PAIRS_PER_PASS = RERANK_THREADS / 4;  // 64 pairs = 128 candidates per pass

__device__ __forceinline__ int base_class(unsigned char c)
{
    switch (c) {
    case 'A': return 0;
    case 'C': return 1;
    case 'G': return 2;
    case 'T': return 3;
    case 'N': return 4;
    default: return 7;
    }
}

__global__ void classify_kernel(const uint8_t *__restrict__ raw, uint8_t *__restrict__ cls, long long n, int *has_other)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long step = (long long)gridDim.x * blockDim.x;
    int other = 0;
    for (; i < n; i += step) {
        int c = base_class(raw[i]);
        cls[i] = (uint8_t)c;
        other |= c == 7;
    }
    if (__any_sync(FULL, other) && (threadIdx.x & 31) == 0) atomicExch(has_other, 1);
}

// ---- candidate expansion: find_sequences + the re-expansion in post_process_sw_* (src/utils/post_processor.cpp
// :96-102,117-125,215-236,249-257,404-413,501-510), in size_t arithmetic like the reference.
struct ExpandRule { int mode; u64 stride, bound; };
__device__ __forceinline__ void expand_one(const ExpandRule &e, u64 s, u64 &first, int &count)
{
    if (e.stride == 1) {
        first = s;
        count = (e.mode == 1 || s < e.bound) ? 1 : 0;
        return;
    }
    u64 p = s * e.stride;
    if (p >= e.bound) { first = 0; count = 0; return; }
    u64 start = p >= e.stride - 1 ? p - e.stride + 1 : 0;
    u64 end = p + e.stride < e.bound ? p + e.stride : e.bound;
    first = start;
    count = (int)(end - start);
}

struct RerankDev {
    const uint8_t *ref; int form; long long ref_n; int ref_len;  // classes; genome length or window count
    ExpandRule rule;
    const long long *neighbors; long long n, k_in; int take;
    const char *q_flat; const long long *q_off;
    short *scr_scores; u64 *scr_ids; int C;   // scratch for the batch: [n][C]
    int *n_cands; int *min_cands; u64 *cells;
};

// class of base `i` of window `id`, 6 = "never matches" (window overruns / outside the table)
__device__ __forceinline__ int window_base(const RerankDev &p, u64 id, bool valid, int i)
{
    if (!valid) return 6;
    if (p.form == 1) return p.ref[id * (u64)p.ref_len + (u64)i];
    u64 pos = id >> 1;
    if (id & 1) {
        int c = p.ref[pos + (u64)(p.ref_len - 1 - i)];
        return c < 4 ? 3 - c : c;
    }
    return p.ref[pos + (u64)i];
}
__device__ __forceinline__ bool window_valid(const RerankDev &p, u64 id)
{
    if (p.form == 1) return id < (u64)p.ref_n;
    u64 pos = id >> 1;
    return pos + (u64)p.ref_len <= (u64)p.ref_n;  // find_sequence returns "" otherwise
}

template <bool QN>
__device__ __forceinline__ unsigned sw_quad(const uint8_t *selrow, int rows, const unsigned *Qs, const unsigned *QNs, int lane)
{
    const int l = lane & 3;
    unsigned Q[CPL], H[CPL];
#pragma unroll
    for (int j = 0; j < CPL; ++j) { Q[j] = Qs[CPL * l + j]; H[j] = 0; }
    unsigned best = 0, last_out = 0, prev_recv = 0;
    const unsigned NEG1 = 0xffffffffu;
    for (int t = 0; t < rows + 3; ++t) {
        unsigned recv = __shfl_up_sync(FULL, last_out, 1, 4);
        if (l == 0) recv = 0;
        const unsigned x8 = selrow[t - l + 3];
        const unsigned sel = (x8 & 7u) | ((x8 & 0x70u) << 4) | 0x5050u;
        unsigned diag = prev_recv, left = recv, mprev = 0;
#pragma unroll
        for (int j = 0; j < CPL; ++j) {
            const unsigned s2 = __byte_perm(Q[j], QN ? QNs[CPL * l + j] : 0u, sel);
            const unsigned x = diag + s2;
            const unsigned up = H[j];
            const unsigned m = __vimax3_s16x2(x, up, left);
            const unsigned h = __viaddmax_s16x2_relu(m, NEG1, 0u);
            if (j & 1) best = __vimax3_s16x2(best, m, mprev);
            mprev = m;
            diag = up;
            left = h;
            H[j] = h;
        }
        prev_recv = recv;
        last_out = left;
    }
    best = __vmaxs2(best, __shfl_xor_sync(FULL, best, 1));
    best = __vmaxs2(best, __shfl_xor_sync(FULL, best, 2));
    return best;  // per half: max(H) + 1 (or 0)
}

// list == nullptr: one CTA per read (r = blockIdx.x).  Otherwise the CTAs stride over list[0 .. *list_n): the reads the
// profile kernel handed back (those containing N).
__global__ void __launch_bounds__(RERANK_THREADS, 2) sw_rerank_kernel(const RerankDev p, const int *list, const int *list_n)
{
    __shared__ u64 s_cand[MAXC];
    __shared__ u64 s_first[MAXTAKE];
    __shared__ int s_count[MAXTAKE + 1];
    __shared__ unsigned s_Q[QCOLS], s_QN[QCOLS];
    __shared__ __align__(4) uint8_t s_sel[PAIRS_PER_PASS][MAX_FAST_REFLEN + 8];
    __shared__ int s_nc, s_hasN;

    const int tid = threadIdx.x, lane = tid & 31;
    const int n_list = list ? *list_n : 1;
    for (int li = list ? (int)blockIdx.x : 0; li < n_list; li += list ? (int)gridDim.x : 1) {
    const long long r = list ? (long long)list[li] : (long long)blockIdx.x;
    const long long qb = p.q_off[r];
    const int lq = (int)(p.q_off[r + 1] - qb);
    const int rows = p.ref_len;

    // query profile: byte c of s_Q[j] = 2 when query base j is class c (0..3), s_QN[j] byte 0 = 2 for 'N'
    __syncthreads();
    if (tid == 0) s_hasN = 0;
    __syncthreads();
    for (int j = tid; j < QCOLS; j += RERANK_THREADS) {
        unsigned q = 0, qn = 0;
        if (j < lq) {
            int c = base_class((unsigned char)p.q_flat[qb + j]);
            if (c < 4) q = 2u << (8 * c);
            else if (c == 4) { qn = 2u; s_hasN = 1; }
        }
        s_Q[j] = q;
        s_QN[j] = qn;
    }
    // candidate list
    const long long *nb = p.neighbors + r * p.k_in;
    for (int i = tid; i < p.take; i += RERANK_THREADS) {
        u64 first; int count;
        expand_one(p.rule, (u64)nb[i], first, count);
        s_first[i] = first;
        s_count[i] = count;
    }
    __syncthreads();
    if (tid == 0) {
        int acc = 0;
        for (int i = 0; i < p.take; ++i) { int c = s_count[i]; s_count[i] = acc; acc += c; }
        s_count[p.take] = acc;
        s_nc = acc;
    }
    __syncthreads();
    const int nc = s_nc;  // host guarantees take * per <= MAXC
    for (int i = tid; i < p.take; i += RERANK_THREADS) {
        int b = s_count[i], e = s_count[i + 1];
        u64 f = s_first[i];
        for (int c = b; c < e; ++c) s_cand[c] = f + (u64)(c - b);
    }
    if (tid == 0) {
        if (p.n_cands) p.n_cands[r] = nc;
        if (p.min_cands) atomicMin(p.min_cands, nc);
        if (p.cells) atomicAdd(p.cells, (u64)nc * (u64)rows * (u64)lq);
    }
    __syncthreads();

    const bool hasN = s_hasN != 0;
    const int pair = tid >> 2;
    for (int c0 = 0; c0 < nc; c0 += 2 * PAIRS_PER_PASS) {
        // stage the (base_a | base_b << 4) rows of this pass, 3 never-match rows of padding either side
        const int npairs = min(PAIRS_PER_PASS, (nc - c0 + 1) / 2);
        for (int e = tid; e < npairs * (rows + 6); e += RERANK_THREADS) {
            int pr = e / (rows + 6), i = e % (rows + 6) - 3;
            int ca = c0 + 2 * pr, cb = ca + 1;
            unsigned v = 0x66;
            if (i >= 0 && i < rows) {
                u64 ida = s_cand[ca];
                int a = window_base(p, ida, window_valid(p, ida), i);
                int b = 6;
                if (cb < nc) { u64 idb = s_cand[cb]; b = window_base(p, idb, window_valid(p, idb), i); }
                v = (unsigned)a | ((unsigned)b << 4);
            }
            s_sel[pr][i + 3] = (uint8_t)v;
        }
        __syncthreads();
        // warps whose 8 pairs are all beyond npairs skip the DP (warp-uniform)
        if ((tid >> 5) * 8 < npairs) {
            const uint8_t *selrow = s_sel[pair < npairs ? pair : 0];
            unsigned best = hasN ? sw_quad<true>(selrow, rows, s_Q, s_QN, lane) : sw_quad<false>(selrow, rows, s_Q, s_QN, lane);
            if ((lane & 3) == 0 && pair < npairs) {
                int sa = (int)(short)(best & 0xffffu) - 1, sb = (int)(short)(best >> 16) - 1;
                int ca = c0 + 2 * pair;
                size_t base = (size_t)r * p.C;
                p.scr_scores[base + ca] = (short)(sa > 0 ? sa : 0);
                p.scr_ids[base + ca] = s_cand[ca];
                if (ca + 1 < nc) {
                    p.scr_scores[base + ca + 1] = (short)(sb > 0 ? sb : 0);
                    p.scr_ids[base + ca + 1] = s_cand[ca + 1];
                }
            }
        }
        __syncthreads();
    }
    }  // reads of this CTA
}

// ---------------------------------------------------------------------------------------------
// Profile kernel (sw_rerank_profile_kernel): same quad pipeline, but the substitution term comes from a
// per-read QUERY PROFILE in shared memory instead of a PRMT per cell pair, which moves the lookup off the
// ALU pipe (PRMT issues at half the VIMNMX rate on sm_100a, tools/ubench_alu.cu) onto the LSU:
//   tab[ab][g][slot] : uint4 = the four packed terms (0 / 2 per 16-bit half) of query columns
//                      38*(slot&3) + 4g .. +3 against the window-base pair ab = 5a + b (a, b in A,C,G,T,none)
// One LDS.128 feeds four cell pairs.  `slot` = lane & 7 replicates the table for the two quads of a
// quarter-warp, so the eight 16-byte addresses of every LDS.128 wavefront fall in eight different bank
// groups whatever base pairs the lanes are looking up: conflict-free by construction.
// Used when neither the read nor (effectively) the windows contain N: a window N can only match a read N.
constexpr int PCPL = 38;             // query columns per lane: 4 x 38 = 152 = 150 + the '<' '>' decoration of format_fastq
constexpr int PGROUPS = (PCPL + 3) / 4;  // 10 table groups of 4 columns; the last one is half used
constexpr int PQCOLS = 4 * PGROUPS * 4;  // 160: s_qcls is indexed up to 38 * 3 + 39
constexpr int PTAB_ENTRIES = 25 * PGROUPS * 8;

// The recurrence runs on s16x2 integers whose values stay below 2048.  Such an integer k has the same bit pattern as the
// fp16 number k * 2^-24 (a subnormal below 1024, a normal with ulp 2^-24 above), on which fp16 addition is exact, so
// every step exists in two interchangeable forms that produce identical bits:
//     add     x = s + diag      IMAD (fma pipe)              |  HADD2 / HFMA2 (fma pipe)
//     clamp   h = max(m-1, 0)   VIADDMNMX.S16x2.RELU (alu)   |  HFMA2.RELU(m, 1.0, -2^-24) (fma pipe)
//     max3    m = max(x,up,left) VIMNMX3.S16x2 (alu) only
// ALU_EVERY picks the mix so that the alu and fma pipes (each one warp instruction per 2 clocks per SMSP) carry
// equal shares:  1 = every clamp on the alu pipe (integer form throughout), 0 = every clamp on the fma pipe,
// n >= 2 = every n-th clamp on the alu pipe, the rest on the fma pipe.
// POFF: the staged rows of a window pair are ONE array of 16-bit byte offsets into the table ((5a + b) * PGROUPS * 8 * 16,
// precomputed by the staging pass) instead of two class bytes, so the per-row table address is one LDS.U16 + one add.
// UNR: rows per trip of the row loop.  ptxas renames the H registers from row to row (operand bank conflicts) and restores the
// mapping with moves at the end of the trip: 9 moves per row at 2 rows per trip, 5 at 8.
template <int ALU_EVERY, bool POFF = false, int UNR = 2>
__device__ __forceinline__ unsigned sw_quad_profile(const uint8_t *rowA, const uint8_t *rowB, int rows, const uint4 *tab, int lane, unsigned one)
{
    constexpr bool F32ADD = ALU_EVERY >= 100;  // ALU_EVERY = 100 + n: the add runs as an fp32 FADD, clamp mix n as below
    constexpr int AE = ALU_EVERY % 100;
    const int l = lane & 3;
    const uint4 *tslot = tab + (lane & 7);
    unsigned H[PCPL];
#pragma unroll
    for (int j = 0; j < PCPL; ++j) H[j] = 0;
    unsigned best = 0, last_out = 0, prev_recv = 0;
    const unsigned NEG1 = 0xffffffffu;
    const unsigned H_ONE = 0x3c003c00u, H_NEGULP = 0x80018001u;  // (1.0h, 1.0h), (-2^-24, -2^-24)
    const unsigned short *poff = reinterpret_cast<const unsigned short *>(rowA) + (3 - l);
#pragma unroll UNR
    for (int t = 0; t < rows + 3; ++t) {
        unsigned recv = __shfl_up_sync(FULL, last_out, 1, 4);
        if (l == 0) recv = 0;
        const uint4 *prow = POFF ? reinterpret_cast<const uint4 *>(reinterpret_cast<const unsigned char *>(tslot) + poff[t])
                                 : tslot + (5u * rowA[t - l + 3] + rowB[t - l + 3]) * (PGROUPS * 8);
        unsigned diag = prev_recv, left = recv, mprev = 0;
#pragma unroll
        for (int g = 0; g < PGROUPS; ++g) {
            const uint4 pr = prow[g * 8];  // the last group uses two of its four columns (an LDS.64 there would be 2-way conflicted)
            const unsigned s4[4] = {pr.x, pr.y, pr.z, pr.w};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int j = 4 * g + c;
                if (j >= PCPL) continue;
                const unsigned up = H[j];
                unsigned x, h;
                if (F32ADD) {
                    // the packed pair (hi << 16 | lo, both < 256) is below 2^24, and an fp32 whose bit pattern p is below 2^24 has
                    // the value p * 2^-149 (denormal, or exponent field 1 with the same ulp): a non-flushing fp32 add of two such
                    // patterns IS the integer add - and FADD can issue on either fma pipe half, which HADD2 / IMAD cannot
                    float xf;
                    asm("add.rn.f32 %0, %1, %2;" : "=f"(xf) : "f"(__uint_as_float(s4[c])), "f"(__uint_as_float(diag)));
                    x = __float_as_uint(xf);
                } else if (AE == 1) x = s4[c] * one + diag;
                else asm("add.rn.f16x2 %0, %1, %2;" : "=r"(x) : "r"(s4[c]), "r"(diag));
                const unsigned m = __vimax3_s16x2(x, up, left);
                const bool on_alu = AE == 1 || (AE >= 2 && (j % AE) == AE - 1);
                if (on_alu) h = __viaddmax_s16x2_relu(m, NEG1, 0u);
                else asm("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(h) : "r"(m), "r"(H_ONE), "r"(H_NEGULP));
                if (j & 1) best = __vimax3_s16x2(best, m, mprev);
                mprev = m;
                diag = up;
                left = h;
                H[j] = h;
            }
        }
        prev_recv = recv;
        last_out = left;
    }
    best = __vmaxs2(best, __shfl_xor_sync(FULL, best, 1));
    best = __vmaxs2(best, __shfl_xor_sync(FULL, best, 2));
    return best;  // per half: max(H) + 1 (or 0)
}

// Two window rows per iteration, skewed by one column: cell (r0, k) and cell (r0 + 1, k - 1) do not depend on each
// other, so every thread carries TWO independent max3 -> clamp chains and the fixed latency of one hides behind the
// other (the single-row form stalls on its own chain with 4 warps per scheduler: "stall_wait" in
// profiles/r1_ncu_sw_rerank_profile_kernel_blend4.txt).  Row r0 + 1 reads its `up` values straight from the registers
// row r0 has just produced, so H[] is still one register per column.  Row r0's clamps go to the fma pipe; row
// r0 + 1's go to the alu pipe at every AE-th column (AE = 0: never).
template <int AE>
__device__ __forceinline__ unsigned sw_quad_profile2(const uint8_t *rowA, const uint8_t *rowB, int rows, const uint4 *tab, int lane)
{
    const int l = lane & 3;
    const uint4 *tslot = tab + (lane & 7);
    unsigned H[PCPL];
#pragma unroll
    for (int j = 0; j < PCPL; ++j) H[j] = 0;
    unsigned best = 0, out0 = 0, out1 = 0, prev_recv1 = 0;
    const unsigned NEG1 = 0xffffffffu;
    const unsigned H_ONE = 0x3c003c00u, H_NEGULP = 0x80018001u;
    const int iters = (rows + 1) / 2 + 3;
#pragma unroll 2
    for (int u = 0; u < iters; ++u) {
        unsigned recv0 = __shfl_up_sync(FULL, out0, 1, 4);
        unsigned recv1 = __shfl_up_sync(FULL, out1, 1, 4);
        if (l == 0) { recv0 = 0; recv1 = 0; }
        const unsigned a2 = *reinterpret_cast<const unsigned short *>(rowA + 2 * (u - l) + 6);
        const unsigned b2 = *reinterpret_cast<const unsigned short *>(rowB + 2 * (u - l) + 6);
        const uint4 *p0 = tslot + (5u * (a2 & 0xffu) + (b2 & 0xffu)) * (PGROUPS * 8);
        const uint4 *p1 = tslot + (5u * (a2 >> 8) + (b2 >> 8)) * (PGROUPS * 8);
        unsigned diag0 = prev_recv1, left0 = recv0;  // row r0: its diagonal neighbour is row r0 - 1's boundary
        unsigned diag1 = recv0, left1 = recv1;       // row r0 + 1
        unsigned h0_last = 0, m0 = 0;
        uint4 q0 = p0[0], q1 = p1[0];
#pragma unroll
        for (int k = 0; k <= PCPL; ++k) {
            unsigned m1 = 0;
            if (k >= 1) {  // row r0 + 1, column k - 1 (uses row r0's column k - 1 from the previous stage)
                const int j = k - 1;
                const unsigned s1 = (j & 3) == 0 ? q1.x : (j & 3) == 1 ? q1.y : (j & 3) == 2 ? q1.z : q1.w;
                unsigned x1, h1;
                asm("add.rn.f16x2 %0, %1, %2;" : "=r"(x1) : "r"(s1), "r"(diag1));
                const unsigned up1 = h0_last;
                m1 = __vimax3_s16x2(x1, up1, left1);
                if (AE >= 1 && (j % AE) == AE - 1) h1 = __viaddmax_s16x2_relu(m1, NEG1, 0u);
                else asm("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(h1) : "r"(m1), "r"(H_ONE), "r"(H_NEGULP));
                diag1 = up1;
                left1 = h1;
                H[j] = h1;
                if ((j & 3) == 3 && j + 1 < PCPL) q1 = p1[((j + 1) >> 2) * 8];
            }
            if (k < PCPL) {  // row r0, column k
                const unsigned s0 = (k & 3) == 0 ? q0.x : (k & 3) == 1 ? q0.y : (k & 3) == 2 ? q0.z : q0.w;
                unsigned x0, h0;
                asm("add.rn.f16x2 %0, %1, %2;" : "=r"(x0) : "r"(s0), "r"(diag0));
                const unsigned up0 = H[k];
                m0 = __vimax3_s16x2(x0, up0, left0);
                asm("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(h0) : "r"(m0), "r"(H_ONE), "r"(H_NEGULP));
                diag0 = up0;
                left0 = h0;
                h0_last = h0;
                if ((k & 3) == 3 && k + 1 < PCPL) q0 = p0[((k + 1) >> 2) * 8];
                best = __vimax3_s16x2(best, m0, m1);
            } else {
                best = __vmaxs2(best, m1);
            }
        }
        prev_recv1 = recv1;
        out0 = left0;
        out1 = left1;
    }
    best = __vmaxs2(best, __shfl_xor_sync(FULL, best, 1));
    best = __vmaxs2(best, __shfl_xor_sync(FULL, best, 2));
    return best;
}

// Block-wide exclusive scan of s_count[0 .. n) in place (n <= MAXTAKE), total returned to every thread.  THREADS / 32 <= 32.
template <int THREADS> __device__ __forceinline__ int block_exclusive_scan(int *s_count, int n, int *s_warp, int tid)
{
    const int per = (n + THREADS - 1) / THREADS, b0 = tid * per;
    int sum = 0;
    for (int i = b0; i < min(b0 + per, n); ++i) sum += s_count[i];
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(FULL, inc, o);
        if ((tid & 31) >= o) inc += v;
    }
    if ((tid & 31) == 31) s_warp[tid >> 5] = inc;
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) {
        const int v = s_warp[w];
        if (w < (tid >> 5)) base += v;
        total += v;
    }
    int acc = base + inc - sum;
    for (int i = b0; i < min(b0 + per, n); ++i) { const int c = s_count[i]; s_count[i] = acc; acc += c; }
    if (tid == THREADS - 1) s_count[n] = total;
    return total;
}

// Four consecutive rows rr0 .. rr0 + 3 of window `id` as class bytes (0..3; 4 = matches nothing: outside the window, N, or an
// id whose window overruns the reference).  Reads the two aligned words that cover the four source bases (the class array
// carries 16 bytes of padding either side), funnel-shifts them into place, and for a reverse-complement id reverses the
// bytes and complements them (c ^ 3).
__device__ __forceinline__ unsigned stage_classes4(const RerankDev &p, u64 id, int rows, int rr0)
{
    if (!window_valid(p, id)) return 0x04040404u;
    const bool rc = p.form == 0 && (id & 1);
    const uint8_t *src = p.form == 1 ? p.ref + id * (u64)rows : p.ref + (id >> 1);
    const uint8_t *a = rc ? src + (rows - 4 - rr0) : src + rr0;  // lowest source address of the four
    const unsigned long long ai = reinterpret_cast<unsigned long long>(a);
    const unsigned *aw = reinterpret_cast<const unsigned *>(ai & ~3ull);
    const unsigned lo = __ldg(aw), hi = __ldg(aw + 1);
    unsigned v = __funnelshift_r(lo, hi, ((unsigned)ai & 3u) * 8u);
    if (rc) v = __byte_perm(v, 0u, 0x0123) ^ 0x03030303u;
    const unsigned big = ((v >> 2) & 0x01010101u) * 0xffu;  // bytes holding a class >= 4 (N, other)
    v = (v & ~big) | (0x04040404u & big);
    // rows outside [0, rows) match nothing: `lead` bytes at the low end, `trail` at the high end (0 for every interior word), as one
    // branch-free mask - the clamped funnel shifts give 0 for a shift of 32 (this fix-up was a predicated per-byte loop run by
    // every task: 6 % of the kernel's source-attributed instructions)
    const int lead = min(max(-rr0, 0), 4), trail = min(max(rr0 + 4 - rows, 0), 4);
    const unsigned keep = __funnelshift_lc(0u, 0xffffffffu, 8u * (unsigned)lead) & __funnelshift_rc(0xffffffffu, 0u, 8u * (unsigned)trail);
    return (v & keep) | (0x04040404u & ~keep);
}

// Dynamic shared memory layout (bytes): tab | sel[2 * PAIRS][selw] (one byte row per candidate) | cand[C] | first[take] | count[take+1]
// THREADS = 256 (2 CTAs / SM, 128 candidates per pass) or 128 (4 CTAs / SM, 64 per pass).
// DENSE: ask ptxas for 768 instead of 512 resident threads per SM (80 instead of 117 registers: 3 CTAs of 256 per SM).
// POFF (one row per iteration only): sel holds, per window PAIR, one row of 16-bit table offsets (sw_quad_profile) - the same bytes.
template <int RPI, int ALU_EVERY, int THREADS, bool DENSE, bool POFF = false, int UNR = 2>
__global__ void __launch_bounds__(THREADS, (DENSE ? 768 : 512) / THREADS) sw_rerank_profile_kernel(const RerankDev p, int selw, unsigned one, int *fb_list, int *fb_n)
{
    constexpr int PAIRS = THREADS / 4;
    extern __shared__ __align__(16) unsigned char dsm[];
    uint4 *tab = reinterpret_cast<uint4 *>(dsm);
    uint8_t *s_sel = dsm + PTAB_ENTRIES * sizeof(uint4);
    u64 *s_cand = reinterpret_cast<u64 *>(s_sel + (size_t)2 * PAIRS * selw);
    u64 *s_first = s_cand + p.C;
    int *s_count = reinterpret_cast<int *>(s_first + p.take);
    __shared__ unsigned char s_qcls[PQCOLS];
    __shared__ int s_warp[THREADS / 32];
    __shared__ int s_hasN;

    const int tid = threadIdx.x, lane = tid & 31;
    const long long r = blockIdx.x;
    const long long qb = p.q_off[r];
    const int lq = (int)(p.q_off[r + 1] - qb);
    const int rows = p.ref_len;

    if (tid == 0) s_hasN = 0;
    __syncthreads();
    for (int j = tid; j < PQCOLS; j += THREADS) {
        int c = 5;  // beyond the read, or a byte outside ACGT: matches nothing
        if (j < lq) {
            c = base_class((unsigned char)p.q_flat[qb + j]);
            if (c == 4) s_hasN = 1;
        }
        s_qcls[j] = (unsigned char)c;
    }
    const long long *nb = p.neighbors + r * p.k_in;
    for (int i = tid; i < p.take; i += THREADS) {
        u64 first; int count;
        expand_one(p.rule, (u64)nb[i], first, count);
        s_first[i] = first;
        s_count[i] = count;
    }
    __syncthreads();
    if (s_hasN) {  // a read with N: left to the PRMT kernel (second launch over flagged reads)
        if (tid == 0) fb_list[atomicAdd(fb_n, 1)] = (int)r;
        return;
    }
    const int nc = block_exclusive_scan<THREADS>(s_count, p.take, s_warp, tid);  // host guarantees take * per <= MAXC
    // query profile
    for (int e = tid; e < PTAB_ENTRIES; e += THREADS) {
        const int slot = e & 7, g = (e >> 3) % PGROUPS, ab = (e >> 3) / PGROUPS;
        const int a = ab / 5, b = ab % 5, col = PCPL * (slot & 3) + 4 * g;
        unsigned w[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int q = s_qcls[col + c];
            w[c] = (q == a ? 2u : 0u) | (q == b ? 0x20000u : 0u);  // a, b == 4 ("none") never equal q in 0..3 or 5
        }
        tab[e] = make_uint4(w[0], w[1], w[2], w[3]);
    }
    __syncthreads();
    for (int i = tid; i < p.take; i += THREADS) {
        int b = s_count[i], e = s_count[i + 1];
        u64 f = s_first[i];
        for (int c = b; c < e; ++c) s_cand[c] = f + (u64)(c - b);
    }
    if (tid == 0) {
        if (p.n_cands) p.n_cands[r] = nc;
        if (p.min_cands) atomicMin(p.min_cands, nc);
        if (p.cells) atomicAdd(p.cells, (u64)nc * (u64)rows * (u64)lq);
    }
    __syncthreads();

    const int pair = tid >> 2;
    const int wpr = selw >> 2;  // 32-bit words per staged row
    for (int c0 = 0; c0 < nc; c0 += 2 * PAIRS) {
        const int npairs = min(PAIRS, (nc - c0 + 1) / 2);
        constexpr int PAD = 3 * RPI;             // "none" rows either side: pipeline fill and drain of the quad
        // Stage the windows of this pass, four rows per thread-task: s_sel[c][r + PAD] = class of row r (0..3), 4 = matches
        // nothing (outside the window, N, or an id whose window overruns the reference).  A task reads the two aligned words
        // that cover its four source bases (the class array carries 16 bytes of padding either side), funnel-shifts them
        // into place, and for a reverse-complement id reverses the bytes and complements them (c ^ 3).
        if (POFF) {
            // one task = four rows of a PAIR: offsets (5a + b) * 1280 of the pair's table rows, four 16-bit values
            const int ntask = npairs * wpr;
            int c = tid / wpr, w = tid - c * wpr;
            const int dc = THREADS / wpr, dw = THREADS - dc * wpr;
#pragma unroll 2
            for (int e = tid; e < ntask; e += THREADS) {
                const int ca = c0 + 2 * c;
                const int rr0 = 4 * w - PAD;
                const unsigned va = stage_classes4(p, s_cand[ca], rows, rr0);
                const unsigned vb = ca + 1 < nc ? stage_classes4(p, s_cand[ca + 1], rows, rr0) : 0x04040404u;
                const unsigned ab = va * 5u + vb;  // per byte <= 24: no carries
                constexpr unsigned ROWB = PGROUPS * 8 * sizeof(uint4);  // 1280: 24 * 1280 fits 16 bits
                uint2 o;
                o.x = __byte_perm(ab, 0u, 0x4140) * ROWB;
                o.y = __byte_perm(ab, 0u, 0x4342) * ROWB;
                *reinterpret_cast<uint2 *>(s_sel + (size_t)c * 2 * selw + 8 * w) = o;
                c += dc; w += dw;
                if (w >= wpr) { w -= wpr; ++c; }
            }
        } else {
            const int ntask = 2 * npairs * wpr;
            int c = tid / wpr, w = tid - c * wpr;
            const int dc = THREADS / wpr, dw = THREADS - dc * wpr;
#pragma unroll 2
            for (int e = tid; e < ntask; e += THREADS) {
                const int ci = c0 + c;
                unsigned out = 0x04040404u;
                if (ci < nc) {
                    const u64 id = s_cand[ci];
                    if (window_valid(p, id)) {
                        const bool rc = p.form == 0 && (id & 1);
                        const uint8_t *src = p.form == 1 ? p.ref + id * (u64)rows : p.ref + (id >> 1);
                        const int rr0 = 4 * w - PAD;  // row of output byte 0
                        const uint8_t *a = rc ? src + (rows - 4 - rr0) : src + rr0;  // lowest source address of the four
                        const unsigned long long ai = reinterpret_cast<unsigned long long>(a);
                        const unsigned *aw = reinterpret_cast<const unsigned *>(ai & ~3ull);
                        const unsigned lo = __ldg(aw), hi = __ldg(aw + 1);
                        unsigned v = __funnelshift_r(lo, hi, ((unsigned)ai & 3u) * 8u);
                        if (rc) v = __byte_perm(v, 0u, 0x0123) ^ 0x03030303u;
                        const unsigned big = ((v >> 2) & 0x01010101u) * 0xffu;  // bytes holding a class >= 4 (N, other)
                        v = (v & ~big) | (0x04040404u & big);
                        if (rr0 < 0 || rr0 + 4 > rows) {
#pragma unroll
                            for (int b = 0; b < 4; ++b)
                                if ((unsigned)(rr0 + b) >= (unsigned)rows) v = (v & ~(0xffu << (8 * b))) | (4u << (8 * b));
                        }
                        out = v;
                    }
                }
                *reinterpret_cast<unsigned *>(s_sel + (size_t)c * selw + 4 * w) = out;
                c += dc; w += dw;
                if (w >= wpr) { w -= wpr; ++c; }
            }
        }
        __syncthreads();
        if ((tid >> 5) * 8 < npairs) {
            const int pp = pair < npairs ? pair : 0;
            const uint8_t *rowA = s_sel + (size_t)(2 * pp) * selw, *rowB = rowA + selw;
            unsigned best = RPI == 2 ? sw_quad_profile2<ALU_EVERY>(rowA, rowB, rows, tab, lane) : sw_quad_profile<ALU_EVERY, POFF, UNR>(rowA, rowB, rows, tab, lane, one);
            if ((lane & 3) == 0 && pair < npairs) {
                int sa = (int)(short)(best & 0xffffu) - 1, sb = (int)(short)(best >> 16) - 1;
                int ca = c0 + 2 * pair;
                size_t base = (size_t)r * p.C;
                p.scr_scores[base + ca] = (short)(sa > 0 ? sa : 0);
                p.scr_ids[base + ca] = s_cand[ca];
                if (ca + 1 < nc) {
                    p.scr_scores[base + ca + 1] = (short)(sb > 0 ? sb : 0);
                    p.scr_ids[base + ca + 1] = s_cand[ca + 1];
                }
            }
        }
        __syncthreads();
    }
}

// ---- top-k: std::partial_sort(idx, idx+k, idx+n, scores[a] > scores[b]) exactly as libstdc++ runs it
// (heap-select over [0,k) + sort_heap), so equal scores come out in the reference's order
// (src/utils/reranker.cpp:35-48).  One thread per read.
// The heap routines are libstdc++'s __push_heap / __adjust_heap / __heap_select / __sort_heap over an index array, with the
// storage abstracted: thread-local arrays (any candidate count), or shared memory laid out [position][thread] for the dense
// case (<= 128 candidates), where random accesses cost a shared-memory transaction instead of a local-memory one and the
// results leave the block as coalesced rows.
struct SelLocal {
    unsigned short *idx; const short *sc;
    __device__ __forceinline__ unsigned short get(int i) const { return idx[i]; }
    __device__ __forceinline__ void set(int i, unsigned short v) { idx[i] = v; }
    __device__ __forceinline__ bool comp(unsigned short a, unsigned short b) const { return sc[a] > sc[b]; }
};
constexpr int SEL_T = 64, SEL_C = 128;
struct SelShared {
    unsigned short (*idx)[SEL_T]; short (*sc)[SEL_T]; int t;
    __device__ __forceinline__ unsigned short get(int i) const { return idx[i][t]; }
    __device__ __forceinline__ void set(int i, unsigned short v) { idx[i][t] = v; }
    __device__ __forceinline__ bool comp(unsigned short a, unsigned short b) const { return sc[a][t] > sc[b][t]; }
};
template <class A> __device__ __forceinline__ void ps_push_heap(A &h, int hole, int top, unsigned short value)
{
    int parent = (hole - 1) / 2;
    while (hole > top && h.comp(h.get(parent), value)) {
        h.set(hole, h.get(parent));
        hole = parent;
        parent = (hole - 1) / 2;
    }
    h.set(hole, value);
}
template <class A> __device__ __forceinline__ void ps_adjust_heap(A &h, int hole, int len, unsigned short value)
{
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (h.comp(h.get(child), h.get(child - 1))) child--;
        h.set(hole, h.get(child));
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        h.set(hole, h.get(child - 1));
        hole = child - 1;
    }
    ps_push_heap(h, hole, top, value);
}
// std::partial_sort(idx, idx + k, idx + nc, comp): heap-select, then sort_heap
template <class A> __device__ __forceinline__ void ps_partial_sort(A &h, int k, int nc)
{
    if (k >= 2) {
        int parent = (k - 2) / 2;
        for (;;) {
            const unsigned short value = h.get(parent);
            ps_adjust_heap(h, parent, k, value);
            if (parent == 0) break;
            parent--;
        }
    }
    for (int i = k; i < nc; ++i) {
        if (h.comp(h.get(i), h.get(0))) {
            const unsigned short value = h.get(i);
            h.set(i, h.get(0));
            ps_adjust_heap(h, 0, k, value);
        }
    }
    int last = k;
    while (last > 1) {
        --last;
        const unsigned short value = h.get(last);
        h.set(last, h.get(0));
        ps_adjust_heap(h, 0, last, value);
    }
}

__global__ void sw_select_kernel(const short *__restrict__ scr_scores, const u64 *__restrict__ scr_ids, int C,
                                 const int *__restrict__ n_cands, long long n, int k, int *out_scores, u64 *out_ids)
{
    long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const int nc = n_cands[r];
    if (nc < k || nc == 0) return;  // the host raises "Not enough candidates" for the batch
    unsigned short idx[MAXC];
    short sc[MAXC];
    const short *src = scr_scores + (size_t)r * C;
    for (int i = 0; i < nc; ++i) { idx[i] = (unsigned short)i; sc[i] = src[i]; }
    SelLocal h{idx, sc};
    ps_partial_sort(h, k, nc);
    const u64 *ids = scr_ids + (size_t)r * C;
    for (int i = 0; i < k; ++i) {
        out_scores[r * k + i] = sc[idx[i]];
        out_ids[r * k + i] = ids[idx[i]];
    }
}

// C <= SEL_C: one thread per read again, storage in shared memory, cooperative loads and stores.
__global__ void __launch_bounds__(SEL_T) sw_select_smem_kernel(const short *__restrict__ scr_scores, const u64 *__restrict__ scr_ids, int C,
                                                               const int *__restrict__ n_cands, long long n, int k, int *out_scores, u64 *out_ids)
{
    __shared__ unsigned short s_idx[SEL_C][SEL_T];
    __shared__ short s_sc[SEL_C][SEL_T];
    __shared__ int s_nc[SEL_T];
    const int t = threadIdx.x;
    const long long r0 = (long long)blockIdx.x * SEL_T;
    const int nreads = (int)(n - r0 < SEL_T ? n - r0 : SEL_T);
    if (t < nreads) s_nc[t] = n_cands[r0 + t];
    __syncthreads();
    for (int rr = 0; rr < nreads; ++rr) {  // coalesced: the block reads one read's score row at a time
        const int nc = s_nc[rr];
        const short *src = scr_scores + (size_t)(r0 + rr) * C;
        for (int i = t; i < nc; i += SEL_T) { s_sc[i][rr] = src[i]; s_idx[i][rr] = (unsigned short)i; }
    }
    __syncthreads();
    if (t < nreads) {
        const int nc = s_nc[t];
        if (nc >= k && nc > 0) {
            SelShared h{s_idx, s_sc, t};
            ps_partial_sort(h, k, nc);
        }
    }
    __syncthreads();
    for (int rr = 0; rr < nreads; ++rr) {
        const int nc = s_nc[rr];
        if (nc < k || nc == 0) continue;
        const u64 *ids = scr_ids + (size_t)(r0 + rr) * C;
        for (int i = t; i < k; i += SEL_T) {
            const unsigned short j = s_idx[i][rr];
            out_scores[(r0 + rr) * k + i] = s_sc[j][rr];
            out_ids[(r0 + rr) * k + i] = ids[j];
        }
    }
}



// ---------------------------------------------------------------------------------------------
// Generic scorer: one thread per pair, int32, raw byte equality (calc_sw_score verbatim semantics).
// The rolling row lives in a global scratch laid out [column][thread] so that a warp's accesses to
// one column are coalesced.
__device__ __forceinline__ unsigned char comp_raw(unsigned char c)
{
    switch (c & 0x7f) {  // the reference's table has 128 entries (src/utils/parse_inputs.cpp:5-14)
    case 'A': return 'T';
    case 'T': return 'A';
    case 'C': return 'G';
    case 'G': return 'C';
    case 'N': return 'N';
    default: return 0;
    }
}

struct PairView { const uint8_t *a; int la; bool rc; const uint8_t *b; int lb; long long out; };

struct FlatPairs {  // drm_sw_pairs
    const char *a_flat; const long long *a_off; const char *b_flat; const long long *b_off;
    __device__ bool get(long long i, PairView &v) const
    {
        v.a = reinterpret_cast<const uint8_t *>(a_flat) + a_off[i]; v.la = (int)(a_off[i + 1] - a_off[i]); v.rc = false;
        v.b = reinterpret_cast<const uint8_t *>(b_flat) + b_off[i]; v.lb = (int)(b_off[i + 1] - b_off[i]);
        v.out = i;
        return true;
    }
};

struct RerankPairs {  // fallback rerank: pair = (read, candidate slot)
    const uint8_t *raw; int form; long long ref_n; int ref_len;
    const char *q_flat; const long long *q_off; const u64 *scr_ids; const int *n_cands; int C;
    __device__ bool get(long long i, PairView &v) const
    {
        long long r = i / C;
        int c = (int)(i % C);
        if (c >= n_cands[r]) return false;
        u64 id = scr_ids[i];
        v.b = reinterpret_cast<const uint8_t *>(q_flat) + q_off[r]; v.lb = (int)(q_off[r + 1] - q_off[r]);
        v.out = i; v.rc = false; v.la = ref_len;
        if (form == 1) {
            if (id >= (u64)ref_n) v.la = 0;
            v.a = raw + id * (u64)ref_len;
        } else {
            u64 pos = id >> 1;
            if (pos + (u64)ref_len > (u64)ref_n) v.la = 0;
            v.a = raw + pos; v.rc = (id & 1) != 0;
        }
        return true;
    }
};

template <class Src>
__global__ void sw_generic_kernel(const Src src, long long npairs, int *__restrict__ Hs, short *out16, int *out32)
{
    const long long T = (long long)gridDim.x * blockDim.x;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int *H = Hs + tid;
    for (long long i = tid; i < npairs; i += T) {
        PairView v;
        if (!src.get(i, v)) continue;
        for (int j = 0; j <= v.lb; ++j) H[(long long)j * T] = 0;
        int best = 0;
        for (int ia = 0; ia < v.la; ++ia) {
            const unsigned char ca = v.rc ? comp_raw(v.a[v.la - 1 - ia]) : v.a[ia];
            int diag = 0, left = 0;
            for (int j = 1; j <= v.lb; ++j) {
                const int up = H[(long long)j * T];
                int h = diag + (ca == v.b[j - 1] ? 1 : -1);
                h = max(h, up - 1);
                h = max(h, left - 1);
                h = max(h, 0);
                H[(long long)j * T] = h;
                best = max(best, h);
                diag = up;
                left = h;
            }
        }
        if (out16) out16[v.out] = (short)best;
        if (out32) out32[v.out] = best;
    }
}

// candidate lists for the fallback rerank: one thread per read
__global__ void expand_kernel(const RerankDev p)
{
    long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= p.n) return;
    const long long *nb = p.neighbors + r * p.k_in;
    int nc = 0;
    for (int i = 0; i < p.take; ++i) {
        u64 first; int count;
        expand_one(p.rule, (u64)nb[i], first, count);
        for (int c = 0; c < count; ++c) p.scr_ids[(size_t)r * p.C + nc + c] = first + (u64)c;
        nc += count;
    }
    if (p.n_cands) p.n_cands[r] = nc;
    if (p.min_cands) atomicMin(p.min_cands, nc);
    if (p.cells) atomicAdd(p.cells, (u64)nc * (u64)p.ref_len * (u64)(p.q_off[r + 1] - p.q_off[r]));
}

// Stream-ordered scratch that is returned to the pool on every way out of a launch function (an early DRM_CUDA_TRY return included).
struct AsyncScratch {
    cudaStream_t stream;
    void *p[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int n = 0;
    explicit AsyncScratch(cudaStream_t s) : stream(s) {}
    AsyncScratch(const AsyncScratch &) = delete;
    AsyncScratch &operator=(const AsyncScratch &) = delete;
    template <class T> cudaError_t alloc(T **out, size_t bytes)
    {
        void *q = nullptr;
        const cudaError_t e = cudaMallocAsync(&q, bytes ? bytes : 16, stream);
        if (e == cudaSuccess) p[n++] = q;
        *out = static_cast<T *>(q);
        return e;
    }
    ~AsyncScratch()
    {
        for (int i = 0; i < n; ++i) cudaFreeAsync(p[i], stream);
    }
};

}  // namespace

int launch_classify(const uint8_t *raw, uint8_t *cls, int64_t n, int *has_other_dev, cudaStream_t stream)
{
    if (n <= 0) return DRM_OK;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    classify_kernel<<<(unsigned)blocks, 256, 0, stream>>>(raw, cls, n, has_other_dev);
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

int launch_sw_pairs(const char *a_flat, const int64_t *a_off, const char *b_flat, const int64_t *b_off, int64_t n,
                    int max_len_b, int32_t *scores, cudaStream_t stream)
{
    if (n <= 0) return DRM_OK;
    long long threads = n < 148LL * 1024 ? (n + 127) / 128 * 128 : 148LL * 1024;
    AsyncScratch scratch(stream);
    int *H = nullptr;
    DRM_CUDA_TRY(scratch.alloc(&H, (size_t)threads * (size_t)(max_len_b + 1) * sizeof(int)));
    FlatPairs src{a_flat, reinterpret_cast<const long long *>(a_off), b_flat, reinterpret_cast<const long long *>(b_off)};
    sw_generic_kernel<FlatPairs><<<(unsigned)(threads / 128), 128, 0, stream>>>(src, n, H, nullptr, scores);
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

// One batch of reads.  All pointers are device pointers; q_off holds absolute offsets into q_flat.
int launch_rerank(const DeviceRef &ref, const RerankArgs &a, cudaStream_t stream)
{
    if (a.n <= 0) return DRM_OK;
    if (a.k < 0 || a.k_clusters <= 0 || a.stride <= 0 || a.k_in <= 0) return fail(DRM_EINVAL, "bad rerank arguments");
    if ((uint64_t)a.k > (uint64_t)a.k_clusters * 2u * (uint64_t)a.stride)
        return fail(DRM_EKTOOLARGE, "Final k too large. Ensure k < k_clusters * 2 * stride to have enough candidates.");
    if (ref.form == 1 && a.mode == 1) return fail(DRM_EINVAL, "dynamic mode needs a genome reference (drm_ref_open_genome)");
    const int take = (int)(a.k_clusters < a.k_in ? a.k_clusters : a.k_in);
    const long long per = a.stride == 1 ? 1 : 2 * a.stride - 1;
    const long long C = (long long)take * per;
    if (C > MAXC || take > MAXTAKE) return fail(DRM_EUNSUPPORTED, "more than 1024 candidates per read (k_clusters * (2*stride-1))");
    if (a.k > C) return fail(DRM_ENOTENOUGH, "Not enough candidates (" + std::to_string(C) + " < " + std::to_string(a.k) + ")");

    RerankDev p{};
    p.form = ref.form; p.ref_n = ref.len; p.ref_len = ref.ref_len;
    p.rule.mode = a.mode; p.rule.stride = (u64)a.stride;
    if (a.mode == 1) p.rule.bound = (u64)ref.len;                       // genome length in bases
    else if (ref.form == 1) p.rule.bound = (u64)ref.len;                // number of windows
    else p.rule.bound = ref.len >= ref.ref_len ? 2ull * (u64)(ref.len - ref.ref_len + 1) : 0ull;
    p.k_in = a.k_in; p.take = take;
    p.q_flat = a.q_flat;
    p.C = (int)C;
    p.min_cands = a.min_cands; p.cells = reinterpret_cast<u64 *>(a.cells);

    const bool fast = !ref.has_other && a.max_qlen <= QCOLS && ref.ref_len <= MAX_FAST_REFLEN;
    const long long batch_max = 131072;
    short *scr_scores = nullptr; u64 *scr_ids = nullptr; int *ncand_tmp = nullptr; int *H = nullptr;
    const long long B = a.n < batch_max ? a.n : batch_max;
    AsyncScratch scratch(stream);  // freed (stream-ordered) on every return below
    DRM_CUDA_TRY(scratch.alloc(&scr_scores, (size_t)B * C * sizeof(short)));
    DRM_CUDA_TRY(scratch.alloc(&scr_ids, (size_t)B * C * sizeof(u64)));
    if (!a.n_cands) DRM_CUDA_TRY(scratch.alloc(&ncand_tmp, (size_t)a.n * sizeof(int)));
    int *n_cands = a.n_cands ? a.n_cands : ncand_tmp;
    const long long gthreads = 148LL * 1024;
    if (!fast) DRM_CUDA_TRY(scratch.alloc(&H, (size_t)gthreads * (size_t)(a.max_qlen + 1) * sizeof(int)));
    // Kernel variant: 0 = PRMT lookup; query-profile kernels, one row per iteration: 1 = clamps on the alu
    // pipe (integer form), 2 = clamps on the fma pipe (fp16 form), 3 / 4 = every 2nd / 4th clamp on the alu pipe; two rows
    // per iteration: 10 / 11 / 12 / 14 = none / every / every 2nd / every 4th clamp of the second row on the alu pipe.
    // DRM_SW_KERNEL overrides the default, for A/B measurements and the parity tests.
    static const struct { const char *name; int v; } kVariants[] = {{"prmt", 0}, {"profile", 1}, {"profile16", 2}, {"blend2", 3},
        {"blend4", 4}, {"blend4o", 5}, {"blend4o4", 6}, {"blend4o8", 7}, {"blend4u8", 8}, {"blend3o8", 9}, {"r2a0", 10}, {"r2a1", 11}, {"r2a2", 12}, {"r2a4", 14}, {"f32p16", 20}, {"f32b2", 21}, {"f32b4", 22}, {"f32b3", 23}};
    int variant = 9;  // blend3o8 + the 80-register build: profiles/r1_ab_v7_sw_dense.txt, profiles/r2_ab_sw_rowloop.txt
    if (const char *e = getenv("DRM_SW_KERNEL"))
        for (const auto &kv : kVariants)
            if (!strcmp(e, kv.name)) variant = kv.v;
    const int selw = (ref.ref_len + 14 + 3) / 4 * 4;  // rows + fill/drain padding of the two-row form (12) + 2
    // CTA size of the profile kernel: 256 threads (128 candidates per pass, 2 CTAs/SM) or 128 (64 per pass, 4 CTAs/SM, so
    // that more reads are in different phases on one SM).  DRM_SW_THREADS overrides, for A/B measurements.
    int pthreads = 256;
    if (const char *e = getenv("DRM_SW_THREADS")) pthreads = atoi(e) == 128 ? 128 : 256;
    bool pdense = true;  // the 80-register build, 3 CTAs per SM; DRM_SW_DENSE=0 selects the 117-register build (A/B)
    if (const char *e = getenv("DRM_SW_DENSE")) pdense = atoi(e) != 0;
    const size_t psmem = (size_t)PTAB_ENTRIES * sizeof(uint4) + (size_t)(pthreads / 2) * selw + (size_t)C * sizeof(u64) +
                         (size_t)take * sizeof(u64) + (size_t)(take + 1) * sizeof(int) + 16;
    int *fb = nullptr;  // [0] = count, [1..] = reads handed back by the profile kernel
    if (fast && variant != 0) {
        DRM_CUDA_TRY(scratch.alloc(&fb, (size_t)(B + 1) * sizeof(int)));
    }
    if (a.k > 0) {
        DRM_CUDA_TRY(cudaMemsetAsync(a.out_scores, 0, (size_t)a.n * a.k * sizeof(int32_t), stream));
        DRM_CUDA_TRY(cudaMemsetAsync(a.out_ids, 0, (size_t)a.n * a.k * sizeof(uint64_t), stream));
    }
    for (long long b0 = 0; b0 < a.n; b0 += B) {
        const long long nb = a.n - b0 < B ? a.n - b0 : B;
        p.n = nb;
        p.neighbors = reinterpret_cast<const long long *>(a.neighbors) + b0 * a.k_in;
        p.q_off = reinterpret_cast<const long long *>(a.q_off) + b0;
        p.scr_scores = scr_scores; p.scr_ids = scr_ids;
        p.n_cands = n_cands + b0;
        if (fast && variant != 0) {
            p.ref = ref.data;
            DRM_CUDA_TRY(cudaMemsetAsync(fb, 0, sizeof(int), stream));
#define DRM_PROFILE_LAUNCH_T(R, A, T, DN, PO, UN)                                                                                               \
    do {                                                                                                                                       \
        DRM_CUDA_TRY(cudaFuncSetAttribute(sw_rerank_profile_kernel<R, A, T, DN, PO, UN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)psmem)); \
        sw_rerank_profile_kernel<R, A, T, DN, PO, UN><<<(unsigned)nb, T, psmem, stream>>>(p, selw, 1u, fb + 1, fb);                                \
    } while (0)
#define DRM_PROFILE_LAUNCH(R, A, PO, UN)                                        \
    do {                                                                        \
        if (pthreads == 128) DRM_PROFILE_LAUNCH_T(R, A, 128, false, PO, UN);       \
        else if (pdense) DRM_PROFILE_LAUNCH_T(R, A, 256, true, PO, UN);            \
        else DRM_PROFILE_LAUNCH_T(R, A, 256, false, PO, UN);                       \
    } while (0)
            switch (variant) {
            case 1: DRM_PROFILE_LAUNCH(1, 1, false, 2); break;
            case 2: DRM_PROFILE_LAUNCH(1, 0, false, 2); break;
            case 3: DRM_PROFILE_LAUNCH(1, 2, false, 2); break;
            case 4: DRM_PROFILE_LAUNCH(1, 4, false, 2); break;
            case 5: DRM_PROFILE_LAUNCH(1, 4, true, 2); break;
            case 6: DRM_PROFILE_LAUNCH(1, 4, true, 4); break;
            case 7: DRM_PROFILE_LAUNCH(1, 4, true, 8); break;
            case 8: DRM_PROFILE_LAUNCH(1, 4, false, 8); break;
            case 9: DRM_PROFILE_LAUNCH(1, 3, true, 8); break;
            case 20: DRM_PROFILE_LAUNCH(1, 100, false, 2); break;
            case 21: DRM_PROFILE_LAUNCH(1, 102, false, 2); break;
            case 22: DRM_PROFILE_LAUNCH(1, 104, false, 2); break;
            case 23: DRM_PROFILE_LAUNCH(1, 103, false, 2); break;
            case 10: DRM_PROFILE_LAUNCH(2, 0, false, 2); break;
            case 11: DRM_PROFILE_LAUNCH(2, 1, false, 2); break;
            case 12: DRM_PROFILE_LAUNCH(2, 2, false, 2); break;
            default: DRM_PROFILE_LAUNCH(2, 4, false, 2); break;
            }
#undef DRM_PROFILE_LAUNCH_T
#undef DRM_PROFILE_LAUNCH
            DRM_CUDA_TRY(cudaGetLastError());
            // reads containing N were handed back: the PRMT kernel strides over that list (usually empty)
            sw_rerank_kernel<<<(unsigned)(nb < 296 ? nb : 296), RERANK_THREADS, 0, stream>>>(p, fb + 1, fb);
            DRM_CUDA_TRY(cudaGetLastError());
        } else if (fast) {
            p.ref = ref.data;
            sw_rerank_kernel<<<(unsigned)nb, RERANK_THREADS, 0, stream>>>(p, nullptr, nullptr);
            DRM_CUDA_TRY(cudaGetLastError());
        } else {
            expand_kernel<<<(unsigned)((nb + 127) / 128), 128, 0, stream>>>(p);
            DRM_CUDA_TRY(cudaGetLastError());
            RerankPairs src{ref.raw, ref.form, ref.len, ref.ref_len, a.q_flat, p.q_off, scr_ids, p.n_cands, (int)C};
            sw_generic_kernel<RerankPairs><<<(unsigned)(gthreads / 128), 128, 0, stream>>>(src, nb * C, H, scr_scores, nullptr);
            DRM_CUDA_TRY(cudaGetLastError());
        }
        if (a.k > 0) {
            static const bool select_local = getenv("DRM_SW_SELECT") && !strcmp(getenv("DRM_SW_SELECT"), "local");  // A/B
            if (C <= SEL_C && !select_local)
                sw_select_smem_kernel<<<(unsigned)((nb + SEL_T - 1) / SEL_T), SEL_T, 0, stream>>>(scr_scores, scr_ids, (int)C, p.n_cands, nb, (int)a.k,
                                                                                              a.out_scores + b0 * a.k, reinterpret_cast<u64 *>(a.out_ids) + b0 * a.k);
            else
                sw_select_kernel<<<(unsigned)((nb + 63) / 64), 64, 0, stream>>>(scr_scores, scr_ids, (int)C, p.n_cands, nb, (int)a.k,
                                                                               a.out_scores + b0 * a.k, reinterpret_cast<u64 *>(a.out_ids) + b0 * a.k);
            DRM_CUDA_TRY(cudaGetLastError());
        }
    }
    return DRM_OK;
}

}  // namespace drm

// ---------------------------------------------------------------------------------------------------------------------
// drm_sw_pairs_device_noalloc: the generic SW pair kernel above for callers that may not allocate (CUDA graph capture,
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The kernel's DP rows (launch_sw_pairs takes them from cudaMallocAsync) live in ONE file-scope device array
// of DRM_SW_PAIRS_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until exit.  All calls on a device
// share it: calls that overlap on the GPU (different streams) race.  Nothing in it has to start at zero - every thread
// clears its own DP row before each pair - so stale contents of the previous call are harmless.
namespace drm {

namespace {

constexpr int NOALLOC_MAX_LEN_B = 4095;                                   // DRM_SW_PAIRS_NOALLOC_MAX_LEN_B
constexpr long long NOALLOC_INTS = 4ll << 20;                             // 16 MiB = DRM_SW_PAIRS_NOALLOC_SCRATCH_BYTES
static __device__ __align__(256) int g_pairs_scratch[NOALLOC_INTS];

// In host code `g_pairs_scratch` is only a stand-in for the symbol; the device address comes from cudaGetSymbolAddress,
// once per device, with that device current.
constexpr int NOALLOC_MAX_DEV = 64;
int *g_pairs_scratch_dev[NOALLOC_MAX_DEV] = {nullptr};
std::once_flag g_pairs_scratch_once[NOALLOC_MAX_DEV];

// pairs whose b is longer than the caller's max_len_b would overrun their DP row: they are presented to the kernel as empty
// (score 0) and marked -1 afterwards by mark_long_pairs_kernel (SW scores are never negative)
struct FlatPairsBounded {
    const char *a_flat; const long long *a_off; const char *b_flat; const long long *b_off; int max_len_b;
    __device__ bool get(long long i, PairView &v) const
    {
        v.a = reinterpret_cast<const uint8_t *>(a_flat) + a_off[i]; v.la = (int)(a_off[i + 1] - a_off[i]); v.rc = false;
        v.b = reinterpret_cast<const uint8_t *>(b_flat) + b_off[i];
        const long long lb = b_off[i + 1] - b_off[i];
        v.lb = (int)lb;
        if (lb > max_len_b) { v.la = 0; v.lb = 0; }
        v.out = i;
        return true;
    }
};

__global__ void mark_long_pairs_kernel(const long long *__restrict__ b_off, long long n, int max_len_b, int *__restrict__ scores)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && b_off[i + 1] - b_off[i] > max_len_b) scores[i] = -1;
}

}  // namespace

int launch_sw_pairs_noalloc(int device, const char *a_flat, const int64_t *a_off, const char *b_flat, const int64_t *b_off, int64_t n,
                            int max_len_b, int32_t *scores, cudaStream_t stream)
{
    if (device < 0 || device >= NOALLOC_MAX_DEV) return fail(DRM_EINVAL, "drm_sw_pairs_device_noalloc: bad device");
    if (n < 0 || max_len_b < 0 || (n > 0 && (!a_flat || !a_off || !b_flat || !b_off || !scores)))
        return fail(DRM_EINVAL, "bad arguments to drm_sw_pairs_device_noalloc");
    if (max_len_b > NOALLOC_MAX_LEN_B)
        return fail(DRM_EUNSUPPORTED, "drm_sw_pairs_device_noalloc: max_len_b " + std::to_string(max_len_b) + " is above the limit of " +
                                          std::to_string(NOALLOC_MAX_LEN_B) + " (static scratch of 16 MiB); use drm_sw_pairs");
    if (n == 0) return DRM_OK;
    DRM_CUDA_TRY(cudaSetDevice(device));
    cudaError_t serr = cudaSuccess;
    std::call_once(g_pairs_scratch_once[device], [&] {
        int *ptr = nullptr;
        serr = cudaGetSymbolAddress(reinterpret_cast<void **>(&ptr), g_pairs_scratch);
        if (serr == cudaSuccess) g_pairs_scratch_dev[device] = ptr;
    });
    if (serr != cudaSuccess) return fail(DRM_ECUDA, std::string("cudaGetSymbolAddress(g_pairs_scratch): ") + cudaGetErrorString(serr));
    int *H = g_pairs_scratch_dev[device];
    if (!H) return fail(DRM_ECUDA, "drm_sw_pairs_device_noalloc: the static scratch has no device address on this GPU");
    // as many DP rows as fit: one per thread, max_len_b + 1 ints each (at least 1024 threads at the length limit)
    long long threads = NOALLOC_INTS / (max_len_b + 1) / 128 * 128;
    if (threads > 148LL * 1024) threads = 148LL * 1024;
    if (n < threads) threads = (n + 127) / 128 * 128;
    FlatPairsBounded src{a_flat, reinterpret_cast<const long long *>(a_off), b_flat, reinterpret_cast<const long long *>(b_off), max_len_b};
    sw_generic_kernel<FlatPairsBounded><<<(unsigned)(threads / 128), 128, 0, stream>>>(src, n, H, nullptr, scores);
    DRM_CUDA_TRY(cudaGetLastError());
    mark_long_pairs_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(reinterpret_cast<const long long *>(b_off), n, max_len_b, scores);
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

}  // namespace drm

extern "C" int drm_sw_pairs_device_noalloc(int device, const char *a_flat_dev, const int64_t *a_off_dev, const char *b_flat_dev,
                                           const int64_t *b_off_dev, int64_t n, int max_len_b, int32_t *scores_dev, void *stream)
{
    return drm::launch_sw_pairs_noalloc(device, a_flat_dev, a_off_dev, b_flat_dev, b_off_dev, n, max_len_b, scores_dev,
                                        static_cast<cudaStream_t>(stream));
}
