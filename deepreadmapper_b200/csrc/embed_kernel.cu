// The embedder on the device: sequence bytes -> 3-mer tokens -> token embedding -> 2-layer
// bidirectional GRU (hidden 64) -> 128-d vector = [layer-2 forward final state | layer-2 backward final state].
// Replaces Vectorizer::vectorize (src/inference/vectorize.cpp:34-141: Preprocessor::preprocess src/inference/preprocess.cpp:20-42,
// prepareBatch :345-366, OpenVINO inference of models/finetuned_sgn33-new-a-Apr6.xml src/inference/fast_model.cpp:3-224).
// CPU restatement: oracle/embed_oracle.py.  Two forms of the recurrences: the tensor-core form further down (tcgen05.mma + TMEM,
// gru_tc_kernel; the default) and the fp32 CUDA-core kernels described next (DRM_EMBED_TC=0, and weights that are not fp16 numbers).
//
// Work per sequence (T = 123 steps): layer-1 recurrence 2 x T x 192 x 64 FMA, layer-2 input projection
// 2 x T x 192 x 128, layer-2 recurrence 2 x T x 192 x 64  = 12.1 M FMA; the layer-1 input projection is a table
// (only 97 token ids can come out of the tokeniser, so W1 . emb[id] + b is precomputed per id at load).
//   gru_kernel<1>  one CTA = 16 sequences, 384 threads = (direction, gate row); a thread keeps its row of R in
//                  registers, the 16 hidden states sit in shared memory (broadcast float4 reads); after the
//                  matrix-vector products a second phase applies the gates.  Writes the states Y1.
//   proj2_kernel   X2 = W2 . Y1 + b for every (sequence, step); a thread keeps its row of W2 (128 floats).
//   gru_kernel<2>  same recurrence with the input projection read from X2; writes the two final states.
// Scratch (Y1: 63 KB, X2: 189 KB per sequence) is sized for a pass of 148 x 16 x 4 sequences (4 CTA waves).
#include "common.h"

#ifndef DRM_EMU
#include <cuda_fp16.h>
#endif

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace drm {

namespace {

constexpr int EH = 64;        // hidden size
constexpr int EG = 192;       // gate rows per direction (z, r, h)
constexpr int ES = 16;        // sequences per CTA
constexpr int ET_MAX = 128;   // max_len bound
constexpr int ETHREADS = 2 * EG;
constexpr int NSLOT = 97;     // pad + the 96 tokens of src/utils/tok2index.cpp
constexpr int CODE_LT = 8, CODE_GT = 9;

struct SrcDev {
    int mode;
    const char *flat; const long long *off;
    const uint8_t *cls; int form; long long ref_n; int ref_len;
    unsigned long long first_id, id_step;
    const unsigned long long *ids; int decorate;
};

// char2Val after tolower (includes/inference/preprocess.hpp:9-25) with '<' and '>' kept apart
__host__ __device__ __forceinline__ int byte_code(unsigned char c)
{
    switch (c | 0x20) {
    case 'a': return 0;
    case 'c': return 1;
    case 'g': return 2;
    case 't': return 3;
    default: return c == '<' ? CODE_LT : (c == '>' ? CODE_GT : 7);
    }
}

struct SeqView {
    const unsigned char *bytes;  // mode 0
    const uint8_t *win; bool rc; int ref_len; int dec;  // mode 1 (dec = 1: '<' and '>' around the window)
    int n;
};

__host__ __device__ __forceinline__ SeqView seq_view(const SrcDev &s, long long i)
{
    SeqView v;
    v.bytes = nullptr; v.win = nullptr; v.rc = false; v.ref_len = s.ref_len; v.dec = s.decorate ? 1 : 0; v.n = 0;
    if (s.mode == 0) {
        v.bytes = reinterpret_cast<const unsigned char *>(s.flat) + s.off[i];
        v.n = (int)(s.off[i + 1] - s.off[i]);
    } else {  // find_sequence (src/utils/post_processor.cpp:47-66) wrapped in PREFIX / POSTFIX (src/utils/parse_inputs.cpp:340-343)
        const unsigned long long id = s.ids ? s.ids[i] : s.first_id + (unsigned long long)i * s.id_step;
        if (s.form == 1) {
            if (id < (unsigned long long)s.ref_n) { v.win = s.cls + id * (unsigned long long)s.ref_len; v.n = s.ref_len + 2 * v.dec; }
        } else {
            const unsigned long long pos = id >> 1;
            if (pos + (unsigned long long)s.ref_len <= (unsigned long long)s.ref_n) {
                v.win = s.cls + pos; v.rc = (id & 1) != 0; v.n = s.ref_len + 2 * v.dec;
            }
        }
    }
    return v;
}

__host__ __device__ __forceinline__ int seq_code(const SeqView &v, int j)
{
    if (v.bytes) return byte_code(v.bytes[j]);
    if (v.dec) {
        if (j == 0) return CODE_LT;
        if (j == v.n - 1) return CODE_GT;
    }
    const int b = j - v.dec;
    const int c = v.rc ? v.win[v.ref_len - 1 - b] : v.win[b];  // base classes 0..3 = A,C,G,T; 4 = N; 7 = other
    if (c > 3) return 7;
    return v.rc ? 3 - c : c;
}

// hashToken (includes/inference/preprocess.hpp:33-52); returns the table slot + 1, or 0 (the pad id) where the
// reference would read past its 96-entry table.
__host__ __device__ __forceinline__ int hash_slot(int c0, int c1, int c2)
{
    const int v0 = c0 < 4 ? c0 : 7, v1 = c1 < 4 ? c1 : 7, v2 = c2 < 4 ? c2 : 7;
    int h;
    if (c0 == CODE_LT) h = (v1 << 2) + v2;
    else if (c2 == CODE_GT) h = 16 + (v0 << 2) + v1;
    else h = 32 + (v0 << 4) + (v1 << 2) + v2;
    return h < 96 ? h + 1 : 0;
}

// Preprocessor::preprocess (src/inference/preprocess.cpp:20-42): token t of the sequence, 0 beyond its end.
__host__ __device__ __forceinline__ int token_slot(const SeqView &v, int t, int T)
{
    const int ln = v.n < T ? v.n : T;
    if (ln < 2 || t >= ln) return 0;
    if (t == 0) return hash_slot(CODE_LT, seq_code(v, 0), seq_code(v, 1));
    if (t < ln - 1) return hash_slot(seq_code(v, t - 1), seq_code(v, t), seq_code(v, t + 1));
    return hash_slot(seq_code(v, ln - 2), seq_code(v, ln - 1), ln < v.n ? seq_code(v, ln) : CODE_GT);
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// LAYER 1: xp from the token table P1, writes Y1.  LAYER 2: xp from X2, writes the final states to out.
// PREFETCH: the input-projection values of a step's gate phase are loaded into registers before the step's products, so
// the global-memory latency (X2 comes from DRAM in layer 2: `long_scoreboard` 7.2 per issue in profiles/r1_ncu_embed_v1) hides
// behind ~1300 instructions of FMA work.  Same arithmetic in the same order (bit-identical); the default since it was measured (+45 %), DRM_EMBED_PREFETCH=0 turns it off.
template <int LAYER, bool PREFETCH>
__global__ void __launch_bounds__(ETHREADS, 1)
gru_kernel(const SrcDev src, const long long first, const long long n, const int T, const float *__restrict__ P1,
           const float *__restrict__ X2, const float *__restrict__ R, const float *__restrict__ bR, float *__restrict__ Y1,
           float *__restrict__ out)
{
    __shared__ unsigned char tok_s[ES][ET_MAX];
    __shared__ __align__(16) float h_s[2][ES][EH];
    __shared__ float a_s[2][ES][EG];
    const int tid = threadIdx.x, dir = tid / EG, row = tid % EG;
    const long long base = (long long)blockIdx.x * ES;  // first sequence of this CTA inside the pass

    if (LAYER == 1) {
        for (int p = tid; p < ES * T; p += ETHREADS) {
            const int s = p / T, t = p % T;
            int slot = 0;
            if (base + s < n) {
                const SeqView v = seq_view(src, first + base + s);
                slot = token_slot(v, t, T);
            }
            tok_s[s][t] = (unsigned char)slot;
        }
    }
    for (int p = tid; p < 2 * ES * EH; p += ETHREADS) (&h_s[0][0][0])[p] = 0.0f;
    float w[EH];
#pragma unroll
    for (int k = 0; k < EH; ++k) w[k] = R[(size_t)(dir * EG + row) * EH + k];
    __syncthreads();

    constexpr int GATE_ITERS = (ES * EH + EG - 1) / EG;  // (sequence, unit) pairs per thread in the gate phase
    for (int step = 0; step < T; ++step) {
        const int t = dir ? T - 1 - step : step;
        float xz[GATE_ITERS], xr[GATE_ITERS], xh[GATE_ITERS];
        if (PREFETCH) {
#pragma unroll
            for (int m = 0; m < GATE_ITERS; ++m) {
                const int p = row + m * EG;
                xz[m] = xr[m] = xh[m] = 0.0f;
                if (p < ES * EH) {
                    const int s = p / EH, i = p % EH;
                    const long long seq = base + s;
                    const float *xp;
                    if (LAYER == 1) xp = P1 + (size_t)(dir * NSLOT + tok_s[s][t]) * EG;
                    else xp = X2 + ((size_t)(seq < n ? seq : 0) * T + t) * (2 * EG) + dir * EG;
                    xz[m] = xp[i]; xr[m] = xp[EH + i]; xh[m] = xp[2 * EH + i];
                }
            }
        }
        // a = R . h for the 16 sequences, four at a time
#pragma unroll 1
        for (int s = 0; s < ES; s += 4) {
            float acc0 = 0.0f, acc1 = 0.0f, acc2 = 0.0f, acc3 = 0.0f;
            const float4 *h0 = reinterpret_cast<const float4 *>(&h_s[dir][s][0]);
            const float4 *h1 = reinterpret_cast<const float4 *>(&h_s[dir][s + 1][0]);
            const float4 *h2 = reinterpret_cast<const float4 *>(&h_s[dir][s + 2][0]);
            const float4 *h3 = reinterpret_cast<const float4 *>(&h_s[dir][s + 3][0]);
#pragma unroll
            for (int k4 = 0; k4 < EH / 4; ++k4) {
                const float4 a = h0[k4], b = h1[k4], c = h2[k4], d = h3[k4];
                acc0 = fmaf(w[4 * k4], a.x, acc0); acc0 = fmaf(w[4 * k4 + 1], a.y, acc0);
                acc0 = fmaf(w[4 * k4 + 2], a.z, acc0); acc0 = fmaf(w[4 * k4 + 3], a.w, acc0);
                acc1 = fmaf(w[4 * k4], b.x, acc1); acc1 = fmaf(w[4 * k4 + 1], b.y, acc1);
                acc1 = fmaf(w[4 * k4 + 2], b.z, acc1); acc1 = fmaf(w[4 * k4 + 3], b.w, acc1);
                acc2 = fmaf(w[4 * k4], c.x, acc2); acc2 = fmaf(w[4 * k4 + 1], c.y, acc2);
                acc2 = fmaf(w[4 * k4 + 2], c.z, acc2); acc2 = fmaf(w[4 * k4 + 3], c.w, acc2);
                acc3 = fmaf(w[4 * k4], d.x, acc3); acc3 = fmaf(w[4 * k4 + 1], d.y, acc3);
                acc3 = fmaf(w[4 * k4 + 2], d.z, acc3); acc3 = fmaf(w[4 * k4 + 3], d.w, acc3);
            }
            a_s[dir][s][row] = acc0; a_s[dir][s + 1][row] = acc1; a_s[dir][s + 2][row] = acc2; a_s[dir][s + 3][row] = acc3;
        }
        __syncthreads();
        // gates: z = s(xz + az), r = s(xr + ar), g = tanh(xh + r (ah + bRh)), h' = (1 - z) g + z h
        if (!PREFETCH) {
            for (int p = row; p < ES * EH; p += EG) {
                const int s = p / EH, i = p % EH;
                const long long seq = base + s;
                const float *xp;
                if (LAYER == 1) xp = P1 + (size_t)(dir * NSLOT + tok_s[s][t]) * EG;
                else xp = X2 + ((size_t)(seq < n ? seq : 0) * T + t) * (2 * EG) + dir * EG;
                const float z = sigmoidf_(xp[i] + a_s[dir][s][i]);
                const float r = sigmoidf_(xp[EH + i] + a_s[dir][s][EH + i]);
                const float g = tanhf(xp[2 * EH + i] + r * (a_s[dir][s][2 * EH + i] + bR[dir * EH + i]));
                const float hn = (1.0f - z) * g + z * h_s[dir][s][i];
                h_s[dir][s][i] = hn;
                if (LAYER == 1 && seq < n) Y1[((size_t)seq * T + t) * (2 * EH) + dir * EH + i] = hn;
            }
        } else {
#pragma unroll
            for (int m = 0; m < GATE_ITERS; ++m) {
                const int p = row + m * EG;
                if (p < ES * EH) {
                    const int s = p / EH, i = p % EH;
                    const long long seq = base + s;
                    const float z = sigmoidf_(xz[m] + a_s[dir][s][i]);
                    const float r = sigmoidf_(xr[m] + a_s[dir][s][EH + i]);
                    const float g = tanhf(xh[m] + r * (a_s[dir][s][2 * EH + i] + bR[dir * EH + i]));
                    const float hn = (1.0f - z) * g + z * h_s[dir][s][i];
                    h_s[dir][s][i] = hn;
                    if (LAYER == 1 && seq < n) Y1[((size_t)seq * T + t) * (2 * EH) + dir * EH + i] = hn;
                }
            }
        }
        __syncthreads();
    }
    if (LAYER == 2) {
        for (int p = row; p < ES * EH; p += EG) {
            const int s = p / EH, i = p % EH;
            if (base + s < n) out[(size_t)(base + s) * (2 * EH) + dir * EH + i] = h_s[dir][s][i];
        }
    }
}

// X2[row][col] = b2[col] + W2[col] . Y1[row], rows = (sequence, step) of the pass; one CTA = the rows of 16 sequences.
constexpr int PROJ_TILE = 32;
__global__ void __launch_bounds__(ETHREADS, 1)
proj2_kernel(const long long n, const int T, const float *__restrict__ W2, const float *__restrict__ b2,
             const float *__restrict__ Y1, float *__restrict__ X2)
{
    __shared__ __align__(16) float y_s[PROJ_TILE][2 * EH];
    const int col = threadIdx.x;
    float w[2 * EH];
#pragma unroll
    for (int k = 0; k < 2 * EH; ++k) w[k] = W2[(size_t)col * (2 * EH) + k];
    const float bias = b2[col];
    const long long total = n * T;
    const long long begin = (long long)blockIdx.x * ES * T;
    long long end = begin + (long long)ES * T;
    if (end > total) end = total;
    for (long long r0 = begin; r0 < end; r0 += PROJ_TILE) {
        const int nr = (int)((end - r0) < PROJ_TILE ? (end - r0) : PROJ_TILE);
        __syncthreads();
        const float4 *g = reinterpret_cast<const float4 *>(Y1 + (size_t)r0 * (2 * EH));
        for (int p = threadIdx.x; p < nr * (2 * EH / 4); p += ETHREADS) reinterpret_cast<float4 *>(&y_s[0][0])[p] = g[p];
        __syncthreads();
#pragma unroll 1
        for (int r = 0; r < nr; r += 4) {  // rows past nr hold stale values; their sums are not stored
            float acc0 = bias, acc1 = bias, acc2 = bias, acc3 = bias;
            const float4 *y0 = reinterpret_cast<const float4 *>(&y_s[r][0]);
            const float4 *y1 = reinterpret_cast<const float4 *>(&y_s[(r + 1) % PROJ_TILE][0]);
            const float4 *y2 = reinterpret_cast<const float4 *>(&y_s[(r + 2) % PROJ_TILE][0]);
            const float4 *y3 = reinterpret_cast<const float4 *>(&y_s[(r + 3) % PROJ_TILE][0]);
#pragma unroll
            for (int k4 = 0; k4 < 2 * EH / 4; ++k4) {
                const float4 a = y0[k4], b = y1[k4], c = y2[k4], d = y3[k4];
                acc0 = fmaf(w[4 * k4], a.x, acc0); acc0 = fmaf(w[4 * k4 + 1], a.y, acc0);
                acc0 = fmaf(w[4 * k4 + 2], a.z, acc0); acc0 = fmaf(w[4 * k4 + 3], a.w, acc0);
                acc1 = fmaf(w[4 * k4], b.x, acc1); acc1 = fmaf(w[4 * k4 + 1], b.y, acc1);
                acc1 = fmaf(w[4 * k4 + 2], b.z, acc1); acc1 = fmaf(w[4 * k4 + 3], b.w, acc1);
                acc2 = fmaf(w[4 * k4], c.x, acc2); acc2 = fmaf(w[4 * k4 + 1], c.y, acc2);
                acc2 = fmaf(w[4 * k4 + 2], c.z, acc2); acc2 = fmaf(w[4 * k4 + 3], c.w, acc2);
                acc3 = fmaf(w[4 * k4], d.x, acc3); acc3 = fmaf(w[4 * k4 + 1], d.y, acc3);
                acc3 = fmaf(w[4 * k4 + 2], d.z, acc3); acc3 = fmaf(w[4 * k4 + 3], d.w, acc3);
            }
            float *x = X2 + (size_t)(r0 + r) * (2 * EG) + col;
            x[0] = acc0;
            if (r + 1 < nr) x[2 * EG] = acc1;
            if (r + 2 < nr) x[4 * EG] = acc2;
            if (r + 3 < nr) x[6 * EG] = acc3;
        }
    }
}

// vocabulary id of table slot 0..95 - the ids of src/utils/tok2index.cpp:3- written as arithmetic:
// "<xy" -> 7542 + 4x+y, "xy>" -> 7558 + 5(4x+y), "xyz" -> 7559 + 5(4x+y) + z
int slot_token_id(int slot)
{
    if (slot < 16) return 7542 + slot;
    if (slot < 32) return 7558 + 5 * (slot - 16);
    const int s = slot - 32;
    return 7559 + 5 * (s >> 2) + (s & 3);
}

template <class T> int upload(const std::vector<T> &h, T **d)
{
    DRM_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(d), h.size() * sizeof(T)));
    DRM_CUDA_TRY(cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
    return DRM_OK;
}


#ifndef DRM_EMU
// ---------------------------------------------------------------------------------------------------------------------
// Tensor-core form of the recurrences (tcgen05 + TMEM; DRM_EMBED_TC=0 selects the fp32 kernels above).
//
// One CTA = 128 sequences x ONE direction; thread s owns sequence s of the tile (= row s of the accumulator = TMEM lane s).
// Per step the gate pre-activations of all 128 sequences are ONE product on the 5th-generation tensor cores,
//     D[128 seq x gates] (fp32, TMEM)  =  A[128 seq x K] (fp16, shared)  .  B[gates x K]^T (fp16, shared)
// issued by one thread (tcgen05.mma.cta_group::1.kind::f16, M = 128), completion signalled through tcgen05.commit on an
// mbarrier; every thread then reads its row back with tcgen05.ld.32x32b, applies the gates in fp32 and writes its new state
// into the A operand of the next step.
//   * The weights are fp16-exact in the reference's model (IR constants are f16), so B is exact.  The activations are fp32
//     values in (-1, 1): they enter as hi + lo with hi = fp16(h), lo = fp16(h - hi), two products accumulated into the same
//     TMEM columns - a 22-bit significand, absolute error below 2^-24 per term; measured against the fp32 kernels in
//     tests/test_gpu_zz_embed.py.
//   * A = [x_t | h], B = [W | R].  Layer 1: x_t = the token's embedding row (64 fp16 numbers, exact: the table is f16 in the IR),
//     copied from a 97-row copy of the table in shared memory, K = 64 + 64 - the tensor core does W1 . emb[token], the fp32
//     kernels' per-token table P1 and its gather are not used.  Layer 2: x_t = y1_t (hi | lo), K = 128 + 64: the input projection
//     W2 . y1_t is FUSED into the step's product, X2 (189 KB per sequence) is never written.
//   * Accumulator columns of a step: z 0..63 | r 64..127 (W x + R h summed by the tensor core) | the candidate gate's recurrent
//     half 128..191 | its input half 192..255 (apart because r multiplies only the recurrent one).  Issued as few, wide products
//     (N = 192 / 256); see the issue_x / issue_h lambdas.
//   * Operand layout in shared memory (K-major, no swizzle; the canonical UMMA "interleave" layout): 16-byte chunk
//     (row r, k-group g = k / 8) at g * rows * 16 + r * 16  ->  8 rows x 16 B core matrices are contiguous (SBO = 128 B),
//     consecutive k-groups are rows * 16 B apart (LBO).  Thread s writes chunk (s, g): consecutive threads, consecutive
//     16-byte words, no bank conflicts.
//   * Layer 1 hands its states to layer 2 already split into hi | lo halves AND already in the operand layout: Y1h is the A-operand
//     image of every (tile of 128 sequences, step), 64 KB each - the bytes of the fp32 states; layer 2 copies images, it gathers nothing.
constexpr int TC_M = 128;                 // sequences per CTA
constexpr int TC_K1 = EH + EH, TC_K2 = 2 * EH + EH;  // [x_t | h]: layer 1 x = the token embedding (64), layer 2 x = y1_t (128)
constexpr int TC_COLS1 = 256;             // TMEM columns per CTA, layer 1 (192 used; two CTAs per SM)
constexpr int TC_COLS2 = 512;             // layer 2: two accumulator sets of 256 (one CTA per SM)

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

// shared-memory matrix descriptor: start address, leading (k-group) and stride (8-row group) byte offsets in 16-byte units,
// descriptor version 1 (sm_100), no swizzle
__device__ __forceinline__ unsigned long long umma_desc(unsigned saddr, unsigned lbo_bytes, unsigned sbo_bytes)
{
    return (unsigned long long)((saddr >> 4) & 0x3fffu) | ((unsigned long long)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((unsigned long long)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
// instruction descriptor, kind::f16: D = fp32, A = B = fp16, both K-major, dense; N >> 3 at bit 17, M >> 4 at bit 24
__host__ __device__ constexpr unsigned umma_idesc(int m, int n) { return (1u << 4) | ((unsigned)(n >> 3) << 17) | ((unsigned)(m >> 4) << 24); }

__device__ __forceinline__ void umma_f16(unsigned d_tmem, unsigned long long adesc, unsigned long long bdesc, unsigned idesc, unsigned accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_ld16(unsigned taddr, float (&v)[16])
{
    unsigned r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                   "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// The gate nonlinearities on the SFU, flush-to-zero forms (the plain __expf / division lower to range checks and rescaling
// around MUFU that the clamped arguments here never need): 2^x and 1/x to 1-2 ulp.
__device__ __forceinline__ float ex2_ftz(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_ftz(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
constexpr float TC_L2E = 1.4426950408889634f;
// The gate phase is bound by its instruction count (two warps per scheduler: profiles/r2_ab_embed_ws.txt), so every argument
// arrives already in the exponent's units: px = -log2(e) x, pu = -2 log2(e) u, each ONE fma from the accumulator (the biases are
// scaled when they are staged).
// z = s(x), r = s(y) with ONE reciprocal: with A = 1 + e^-x, B = 1 + e^-y:  s(x) = B / (A B), s(y) = A / (A B).  The exponent is capped
// at 2^40 so that the product stays finite (moves a result by less than 1e-12); a large positive argument gives e^-x = 0, s = 1.
__device__ __forceinline__ void gate_zr(float px, float py, float &sx, float &sy)
{
    const float A = 1.0f + ex2_ftz(fminf(px, 40.0f)), B = 1.0f + ex2_ftz(fminf(py, 40.0f));
    const float t = rcp_ftz(A * B);
    sx = t * B;
    sy = t * A;
}
// tanh(u) = 2 / (1 + e^-2u) - 1; no cap: 2^pu = +inf gives 1 / inf = 0 and tanh = -1, 2^pu = 0 gives +1
__device__ __forceinline__ float tc_tanh(float pu) { return fmaf(2.0f, rcp_ftz(1.0f + ex2_ftz(pu)), -1.0f); }

// h -> (hi, lo) fp16 pair, packed two values per word
__device__ __forceinline__ void split2(float a, float b, unsigned &hi, unsigned &lo)
{
    const __half2 h2 = __floats2half2_rn(a, b);
    const float2 f = __half22float2(h2);
    const __half2 l2 = __floats2half2_rn(a - f.x, b - f.y);
    hi = *reinterpret_cast<const unsigned *>(&h2);
    lo = *reinterpret_cast<const unsigned *>(&l2);
}

// Both layers are the same step: gate pre-activations = [x_t | h] . [W | R]^T with K = KX + 64.
//   layer 1: x_t = the token's embedding row (64 fp16 numbers, exact: the table is f16 in the IR), copied from a 97-row copy of
//            the table in shared memory - the recurrence never touches global memory for its input;
//   layer 2: x_t = y1_t (128 values, hi | lo).
// Dynamic shared memory: Bw [K/8][192][16 B] | Ah [K/8][128][16 B] | Al [(layer 1: 8, the state only) K/8][128][16 B] | bias floats |
// embedding rows (layer 1) | mbarrier, TMEM address
template <int LAYER> struct TcLayout {
    static constexpr int KX = LAYER == 1 ? EH : 2 * EH;
    static constexpr int K = KX + EH;
    static constexpr int KG = K / 8, KGX = KX / 8;
    static constexpr int AL_H0 = LAYER == 1 ? 0 : KGX;          // k-group of the state inside Al
    // weights: the input half Bx [KX/8][192][16 B] with its rows in the order h, z, r (one N = 192 product per k-step lands xh | xz | xr on
    // columns 0..191), then the recurrent half Bh [8][192][16 B] with rows z, r, h (one N = 192 product lands on columns 64..255)
    static constexpr int BX_ROWS = EG;
    static constexpr int BX_BYTES = KGX * BX_ROWS * 16, BH_BYTES = (EH / 8) * EG * 16;
    static constexpr int B_BYTES = BX_BYTES + BH_BYTES;
    static constexpr int AH_BYTES = KG * TC_M * 16;
    static constexpr int AL_BYTES = (AL_H0 + EH / 8) * TC_M * 16;
    static constexpr int BIAS_FLOATS = EG + EH;                  // input biases z, r, h | recurrent bias of h
    static constexpr int OFF_AH = B_BYTES, OFF_AL = OFF_AH + AH_BYTES, OFF_BIAS = OFF_AL + AL_BYTES;
    static constexpr int OFF_EMB = OFF_BIAS + BIAS_FLOATS * 4;
    static constexpr int OFF_BAR = OFF_EMB + (LAYER == 1 ? NSLOT * EH * 2 : 0);
    static constexpr int BYTES = OFF_BAR + 16;
};

// TPS = threads per sequence: 1 (128 threads) or 2 (256 threads: warps w and w + 4 read the same 32 TMEM lanes - the hardware
// ties a warp to lanes 32 (w % 4) .. + 31 - and each owns 32 of the 64 hidden units, so the gate phase has twice the warps to hide
// the SFU and TMEM latencies behind).
template <int LAYER, int TPS, bool PROF = false, int XG = 16>
__global__ void __launch_bounds__(TC_M * TPS, LAYER == 1 ? 2 : 1)
gru_tc_kernel(const SrcDev src, const long long first, const long long n, const int T, const uint4 *__restrict__ emb_g,
              const uint4 *__restrict__ Bw_g, const float *__restrict__ bias_g, const __half *__restrict__ Y1h_in,
              __half *__restrict__ Y1h_out, float *__restrict__ out, unsigned long long *__restrict__ phase_clk)
{
    using L = TcLayout<LAYER>;
    constexpr int NT = TC_M * TPS;
    // phase_clk != nullptr (DRM_EMBED_TC_PHASES=1, profiling only): thread 0 of CTA 0 sums the clocks it spends per step in
    // 0 staging + barrier, 1 issuing the products, 2 waiting for them, 3 restaging the next input (layer 2), 4 the gate phase
    unsigned long long pc[5] = {0, 0, 0, 0, 0};
    const bool prof = PROF && phase_clk != nullptr && blockIdx.x == 0 && threadIdx.x == 0;  // (a separate instantiation: the counters cost registers)
    long long tprev = prof ? clock64() : 0;
#define TC_PHASE(i) do { if (prof) { const long long tn = clock64(); pc[i] += (unsigned long long)(tn - tprev); tprev = tn; } } while (0)
    constexpr int UNITS = EH / TPS;  // hidden units per thread
    constexpr int KSX = L::KX / 16, KS = L::K / 16;  // products (K = 16 each) of the input half, of the whole step
    extern __shared__ __align__(1024) unsigned char tsm[];
    uint4 *Bw = reinterpret_cast<uint4 *>(tsm);
    uint4 *Ah = reinterpret_cast<uint4 *>(tsm + L::OFF_AH);
    uint4 *Al = reinterpret_cast<uint4 *>(tsm + L::OFF_AL);
    float *bias = reinterpret_cast<float *>(tsm + L::OFF_BIAS);
    uint4 *embs = reinterpret_cast<uint4 *>(tsm + L::OFF_EMB);     // [slot][8 x 16 B]
    unsigned long long *bar = reinterpret_cast<unsigned long long *>(tsm + L::OFF_BAR);
    unsigned *tmem_slot = reinterpret_cast<unsigned *>(tsm + L::OFF_BAR + 8);

    const int tid = threadIdx.x, warp = tid >> 5;
    const int row = tid & (TC_M - 1), part = tid / TC_M;  // sequence of the tile, which share of its hidden units
    const int dir = blockIdx.x & 1;
    const long long seq = (long long)(blockIdx.x >> 1) * TC_M + row;  // inside the pass
    const bool live = seq < n;

    // ---- one-time set-up: weights of this direction, biases, embedding rows, zero state, mbarrier, TMEM columns
    {
        const uint4 *g = Bw_g + (size_t)dir * (L::B_BYTES / 16);
        for (int p = tid; p < L::B_BYTES / 16; p += NT) Bw[p] = __ldg(g + p);
        // biases z | r | h (input) | h (recurrent), in the units of their exponents (see gate_zr / tc_tanh)
        for (int p = tid; p < L::BIAS_FLOATS; p += NT) bias[p] = bias_g[dir * L::BIAS_FLOATS + p] * (p < 2 * EH ? -TC_L2E : -2.0f * TC_L2E);
        if (LAYER == 1)
            for (int p = tid; p < NSLOT * EH * 2 / 16; p += NT) embs[p] = __ldg(emb_g + p);
        const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
        for (int p = tid; p < (L::AH_BYTES + L::AL_BYTES) / 16; p += NT) Ah[p] = z4;  // Ah and Al are adjacent
    }
    SeqView sv = seq_view(src, first + (live ? seq : 0));  // layer 1: tokens are recomputed per step (3 byte reads + a hash)
    if (!live) sv.n = 0;
    const unsigned bar_s = smem_u32(bar);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    constexpr int TCOLS = LAYER == 1 ? TC_COLS1 : TC_COLS2;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TCOLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = *tmem_slot;
    const unsigned trow = tmem + ((unsigned)((warp & 3) * 32) << 16);  // this warp's 32 lanes

    const unsigned ah_s = smem_u32(Ah), al_s = smem_u32(Al), bw_s = smem_u32(Bw);
    constexpr unsigned A_LBO = TC_M * 16, BX_LBO = L::BX_ROWS * 16, BH_LBO = EG * 16, SBO = 128;
    const unsigned bh_s = bw_s + L::BX_BYTES;
    // Accumulator columns of a set: h input (xh) 0..63 | z 64..127 | r 128..191 | h recurrent 192..255.  The input half's rows are
    // stored h, z, r and the recurrent half's z, r, h, so EACH half is one N = 192 product per k-step onto a contiguous column range
    // (0..191 and 64..255) and the two overlap exactly on z and r, where W x and R h are to be summed.  (A tcgen05.mma of this shape
    // - M = 128, K = 16 - keeps the tensor pipe busy in proportion to N and costs issue / completion latency per instruction: few,
    // wide products, and no padding columns.)  Only the step's first recurrent product is split in two: z, r accumulate onto the
    // input half's result while the h-recurrent columns start from zero.
    constexpr unsigned idesc_192 = umma_idesc(TC_M, 3 * EH), idesc_128 = umma_idesc(TC_M, 2 * EH), idesc_64 = umma_idesc(TC_M, EH);
    constexpr unsigned COL_XH = 0, COL_Z = EH, COL_R = 2 * EH, COL_HN = 3 * EH;
    // input half of a step into accumulator set dcol
    constexpr int NPX = (LAYER == 1 ? 1 : 2) * KSX;  // its products: layer 1's input is exact in fp16 (no lo half)
    auto issue_x_one = [&](unsigned dcol, int p) {   // p is a constant after unrolling
        const int half = p / KSX, ks = p % KSX;
        umma_f16(tmem + dcol + COL_XH, umma_desc((half ? al_s : ah_s) + ks * 2 * A_LBO, A_LBO, SBO), umma_desc(bw_s + ks * 2 * BX_LBO, BX_LBO, SBO),
                 idesc_192, p ? 1u : 0u);
    };
    auto issue_x = [&](unsigned dcol) {
#pragma unroll
        for (int p = 0; p < NPX; ++p) issue_x_one(dcol, p);
    };
    // recurrent half
    auto issue_h = [&](unsigned dcol) {
#pragma unroll
        for (int half = 0; half < 2; ++half)
#pragma unroll
            for (int ks = 0; ks < EH / 16; ++ks) {
                const unsigned a_s = half ? al_s + (L::AL_H0 + 2 * ks) * A_LBO : ah_s + (L::KGX + 2 * ks) * A_LBO;
                const unsigned long long ad = umma_desc(a_s, A_LBO, SBO);
                if (half == 0 && ks == 0) {
                    umma_f16(tmem + dcol + COL_Z, ad, umma_desc(bh_s, BH_LBO, SBO), idesc_128, 1u);
                    umma_f16(tmem + dcol + COL_HN, ad, umma_desc(bh_s + 2 * EH * 16, BH_LBO, SBO), idesc_64, 0u);
                } else {
                    umma_f16(tmem + dcol + COL_Z, ad, umma_desc(bh_s + ks * 2 * BH_LBO, BH_LBO, SBO), idesc_192, 1u);
                }
            }
    };
    // Y1h is laid out as the A-operand IMAGE of every (tile, step): [tile][t][hi | lo][k-group 0..15][row 0..127] 16-byte words,
    // 64 KB per (tile, step).  Layer 1 writes word (g, row) from thread `row`: consecutive threads, consecutive words.  Layer 2 copies
    // the image as it is: thread tid moves words tid, tid + NT, ...: coalesced loads, conflict-free stores, no transposition.
    // (A per-sequence layout made both sides 16-byte gathers across 128 lines per warp instruction: the restaging of the next
    // input was a third of layer 2's step - DRM_EMBED_TC_PHASES=1 prints the clocks per phase.)
    constexpr int YWORDS = 2 * (2 * EH / 8) * TC_M;          // 4096 words per (tile, step): hi image, then lo image
    constexpr int YV = LAYER == 2 ? YWORDS / NT : 1;         // words per thread
    const size_t tile = blockIdx.x >> 1;
    auto y_src = [&](int tt) { return reinterpret_cast<const uint4 *>(Y1h_in) + (tile * T + tt) * YWORDS + tid; };
    auto y_store = [&](const uint4 (&v)[YV]) {
#pragma unroll
        for (int j = 0; j < YV; ++j) {
            const int w = tid + NT * j;  // j is a compile-time constant: hi or lo is decided per unrolled copy
            if (NT * j < YWORDS / 2) Ah[w] = v[j];
            else Al[w - YWORDS / 2] = v[j];
        }
    };
    auto y_prefetch = [&](int tt) {  // 64 KB = 512 lines per (tile, step), shared out over the CTA's threads
        const unsigned char *yp = reinterpret_cast<const unsigned char *>(reinterpret_cast<const uint4 *>(Y1h_in) + (tile * T + tt) * YWORDS);
        constexpr int LP = 512 / NT;  // lines per thread
#pragma unroll
        for (int q = 0; q < LP; ++q) asm volatile("prefetch.global.L2 [%0];" ::"l"(yp + 128 * (tid * LP + q)));
    };
    // layer 1: the embedding row of token (sequence, tt) into the input half of A
    auto x_store = [&](int slot) {
        const uint4 *e = embs + slot * (EH / 8);
#pragma unroll
        for (int g = 0; g < EH / 8 / TPS; ++g) Ah[(g * TPS + part) * TC_M + row] = e[g * TPS + part];
    };

    float h[UNITS];
#pragma unroll
    for (int i = 0; i < UNITS; ++i) h[i] = 0.0f;

    // Layer 2 runs as a two-stage pipeline over two accumulator sets: the INPUT half of step t + 1 (y1_{t+1} . W2: 32 of the 48
    // products, and the only part that reads global memory) is issued while the gates of step t are applied; only the recurrent half
    // (h_t . R2, 16 products) sits between the gates of two consecutive steps.  Layer 1 (two CTAs per SM, one accumulator set each)
    // issues both halves together; the other CTA's gate phase covers them.
    if (LAYER == 2) {
        uint4 v[YV];
        const uint4 *y = y_src(dir ? T - 1 : 0);
#pragma unroll
        for (int g = 0; g < YV; ++g) v[g] = __ldg(y + NT * g);
        y_store(v);
        for (int a = 1; a <= 3 && a < T; ++a) y_prefetch(dir ? T - 1 - a : a);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            issue_x(0u);
        }
    }

    unsigned phase = 0;
    int slot = LAYER == 1 ? token_slot(sv, dir ? T - 1 : 0, T) : 0;
    for (int step = 0; step < T; ++step) {
        const int t = dir ? T - 1 - step : step;
        const unsigned dcol = LAYER == 2 ? (unsigned)(step & 1) * 256u : 0u;
        if (LAYER == 1) x_store(slot);
        uint4 ynext[YV];
        const bool more = LAYER == 2 && step + 1 < T;
        if (more) {  // the next step's input (prefetched into L2 three steps ago) travels while the recurrent products run
            const uint4 *y = y_src(dir ? t - 1 : t + 1);
#pragma unroll
            for (int g = 0; g < YV; ++g) ynext[g] = __ldg(y + NT * g);
            if (step + 4 < T) y_prefetch(dir ? t - 4 : t + 4);  // and the one three steps further leaves DRAM now
        }
        // generic-proxy writes of A (this step's input, last step's state) -> visible to the tensor core's async proxy
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");  // the reads of D below precede the next product
        __syncthreads();
        TC_PHASE(0);
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (LAYER == 1) issue_x(dcol);
            issue_h(dcol);
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_s) : "memory");
        }
        TC_PHASE(1);
        if (LAYER == 1 && step + 1 < T) slot = token_slot(sv, dir ? t - 1 : t + 1, T);  // the next token, while the products run
        {   // wait for the products of this step (the commit covers the input half issued one step earlier too)
            unsigned done = 0, spins = 0;
            do {
                asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                             : "=r"(done) : "r"(bar_s), "r"(phase) : "memory");
                if (!done && ++spins > (1u << 27)) __trap();  // a product that never completes must not hang the GPU
            } while (!done);
            phase ^= 1u;
        }
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        TC_PHASE(2);
        if (LAYER == 2 && more) {  // every earlier product is complete: the input half of A is free for the next step
            y_store(ynext);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (tid == 0) {  // all of its products now (XG = 16, the default); with XG < 16 the rest go between the gates below
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                for (int p = 0; p < XG && p < NPX; ++p) issue_x_one(dcol ^ 256u, p);
            }
        }
        TC_PHASE(3);
        const unsigned tacc = trow + dcol;

        // ---- gates, 16 units at a time: z = s(xz + az), r = s(xr + ar), g = tanh(xh + r (ah + bRh)), h' = (1 - z) g + z h
        // (the z and r columns hold xz + az and xr + ar already summed by the tensor core; biases added here, 21 arithmetic
        // instructions per unit + 3 for the hi | lo split)
        uint4 *yout = nullptr;  // layer 1: this (tile, step)'s image, word (k-group dir * 8, row)
        if (LAYER == 1) yout = reinterpret_cast<uint4 *>(Y1h_out) + (tile * T + t) * YWORDS + (size_t)(dir * (EH / 8)) * TC_M + row;
#pragma unroll
        for (int cc = 0; cc < UNITS / 16; ++cc) {
            const int c = cc + part * (UNITS / 16);  // 16-unit chunk of the hidden state
            float az[16], ar[16], ah[16], xh[16];
            tmem_ld16(tacc + COL_Z + 16 * c, az);
            tmem_ld16(tacc + COL_R + 16 * c, ar);
            tmem_ld16(tacc + COL_HN + 16 * c, ah);
            tmem_ld16(tacc + COL_XH + 16 * c, xh);
            tmem_ld_wait();
            unsigned hi[8], lo[8];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int i = 16 * c + j;
                float z, r;
                gate_zr(fmaf(az[j], -TC_L2E, bias[i]), fmaf(ar[j], -TC_L2E, bias[EH + i]), z, r);
                const float g = tc_tanh(fmaf(r, fmaf(ah[j], -2.0f * TC_L2E, bias[EG + i]), fmaf(xh[j], -2.0f * TC_L2E, bias[2 * EH + i])));
                h[16 * cc + j] = fmaf(z, h[16 * cc + j] - g, g);  // (1 - z) g + z h
                // Layer 2, XG < 16 (A/B only, DRM_EMBED_TC_XG): the next step's input products issued XG at a time every few units
                // instead of all sixteen before the gates, so that warp 0 starts its gates with the others.  Measured SLOWER
                // (12.4 M/s at XG = 2, 12.9 M at 4, 13.15 M at 16; profiles/r2_ab_embed_xg.txt): the issuing thread pays about
                // 100 clocks per product wherever it issues it (the queue in front of the tensor pipe is shallow), and each
                // divergent issue also stops the warp's other 31 lanes.
                constexpr int NG = (NPX + XG - 1) / XG, GSTEP = UNITS / NG < 4 ? UNITS / NG : 4;  // groups; units between two of them
                static_assert(LAYER != 2 || NG == 1 || (GSTEP >= 1 && (NG - 1) * GSTEP <= UNITS), "every group must find its place");
                if (LAYER == 2 && NG > 1 && (16 * cc + j + 1) % GSTEP == 0) {
                    const int gi = (16 * cc + j + 1) / GSTEP;  // group 0 went before the gates
                    if (gi < NG && more && tid == 0) {
#pragma unroll
                        for (int p = gi * XG; p < gi * XG + XG && p < NPX; ++p) issue_x_one(dcol ^ 256u, p);
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) split2(h[16 * cc + 2 * j], h[16 * cc + 2 * j + 1], hi[j], lo[j]);
            Ah[(L::KGX + 2 * c) * TC_M + row] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            Ah[(L::KGX + 2 * c + 1) * TC_M + row] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
            Al[(L::AL_H0 + 2 * c) * TC_M + row] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            Al[(L::AL_H0 + 2 * c + 1) * TC_M + row] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
            if (LAYER == 1) {  // (rows beyond the pass's last sequence are written too: the image is copied whole)
                yout[(2 * c) * TC_M] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                yout[(2 * c + 1) * TC_M] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
                yout[YWORDS / 2 + (2 * c) * TC_M] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                yout[YWORDS / 2 + (2 * c + 1) * TC_M] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
            }
        }
        TC_PHASE(4);
    }
    if (prof) {
#pragma unroll
        for (int i = 0; i < 5; ++i) phase_clk[(LAYER - 1) * 5 + i] = pc[i];
    }
#undef TC_PHASE
    if (LAYER == 2 && live) {
        float4 *o = reinterpret_cast<float4 *>(out + (size_t)seq * (2 * EH) + dir * EH + part * UNITS);
#pragma unroll
        for (int q = 0; q < UNITS / 4; ++q) o[q] = make_float4(h[4 * q], h[4 * q + 1], h[4 * q + 2], h[4 * q + 3]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TCOLS) : "memory");
}

// fp16 image of a weight matrix in the operand layout above: rows x K (row-major fp32, fp16-exact) -> [K/8][192][8 halves] at
// `img`; row r lands at (r + rot) mod 192 (rot = 64 turns the gate order z, r, h into h, z, r)
bool tc_pack(const float *Wm, int rows, int K, int rot, uint16_t *img)
{
    bool exact = true;
    for (int r = 0; r < rows; ++r)
        for (int k = 0; k < K; ++k) {
            const float v = Wm[(size_t)r * K + k];
            const __half hv = __float2half_rn(v);
            if (__half2float(hv) != v) exact = false;
            img[((size_t)(k / 8) * rows + (r + rot) % rows) * 8 + k % 8] = __half_as_ushort(hv);
        }
    return exact;
}
#endif  // DRM_EMU
}  // namespace

// The tokeniser on the host (same code as the kernel's prologue): n x T vocabulary ids, 0-padded.
void tokenize_host(const char *flat, const int64_t *off, int64_t n, int T, uint16_t *tokens)
{
    SrcDev s{};
    s.mode = 0; s.decorate = 1; s.flat = flat; s.off = reinterpret_cast<const long long *>(off);
    for (int64_t i = 0; i < n; ++i) {
        const SeqView v = seq_view(s, i);
        for (int t = 0; t < T; ++t) {
            const int slot = token_slot(v, t, T);
            tokens[(size_t)i * T + t] = (uint16_t)(slot ? slot_token_id(slot - 1) : 0);
        }
    }
}

int embedder_upload(const EmbedWeightsHost &w, int device, int max_len, DeviceEmbedder &e)
{
    if (max_len < 2 || max_len > ET_MAX) return fail(DRM_EINVAL, "embedder max_len must be in [2, 128]");
    if (w.vocab <= slot_token_id(95)) return fail(DRM_EUNSUPPORTED, "embedding table smaller than the tokeniser's largest id (7637)");
    if (w.emb.size() != (size_t)w.vocab * EH || w.W1.size() != 2u * EG * EH || w.R1.size() != 2u * EG * EH || w.B1.size() != 2u * 4 * EH ||
        w.W2.size() != 2u * EG * 2 * EH || w.R2.size() != 2u * EG * EH || w.B2.size() != 2u * 4 * EH)
        return fail(DRM_EINVAL, "embedder weights have unexpected sizes");
    DRM_CUDA_TRY(cudaSetDevice(device));
    e.device = device;
    e.max_len = max_len;
    // layer-1 input projection per token the tokeniser can emit: P1[dir][slot] = W1[dir] . emb[id(slot)] + B1[dir][0:192]
    std::vector<float> P1((size_t)2 * NSLOT * EG), bR1(2 * EH), b2(2 * EG), bR2(2 * EH);
    for (int dir = 0; dir < 2; ++dir) {
        for (int slot = 0; slot < NSLOT; ++slot) {
            const float *x = w.emb.data() + (size_t)(slot == 0 ? 0 : slot_token_id(slot - 1)) * EH;
            for (int row = 0; row < EG; ++row) {
                const float *wr = w.W1.data() + (size_t)(dir * EG + row) * EH;
                float acc = 0.0f;
                for (int k = 0; k < EH; ++k) acc += wr[k] * x[k];
                P1[((size_t)dir * NSLOT + slot) * EG + row] = acc + w.B1[(size_t)dir * 4 * EH + row];
            }
        }
        for (int i = 0; i < EH; ++i) {
            bR1[dir * EH + i] = w.B1[(size_t)dir * 4 * EH + 3 * EH + i];
            bR2[dir * EH + i] = w.B2[(size_t)dir * 4 * EH + 3 * EH + i];
        }
        for (int row = 0; row < EG; ++row) b2[dir * EG + row] = w.B2[(size_t)dir * 4 * EH + row];
    }
    int rc;
    if ((rc = upload(P1, &e.P1)) != DRM_OK) return rc;
    if ((rc = upload(w.R1, &e.R1)) != DRM_OK) return rc;
    if ((rc = upload(bR1, &e.bR1)) != DRM_OK) return rc;
    if ((rc = upload(w.W2, &e.W2)) != DRM_OK) return rc;
    if ((rc = upload(b2, &e.b2)) != DRM_OK) return rc;
    if ((rc = upload(w.R2, &e.R2)) != DRM_OK) return rc;
    if ((rc = upload(bR2, &e.bR2)) != DRM_OK) return rc;
#ifndef DRM_EMU
    {   // operand images for the tensor-core form
        constexpr size_t IMG1 = TcLayout<1>::B_BYTES / 2, IMG2 = TcLayout<2>::B_BYTES / 2;  // halves per direction
        std::vector<uint16_t> b1(2 * IMG1, 0), b2i(2 * IMG2, 0);
        std::vector<float> bias1((size_t)2 * (EG + EH)), bias2((size_t)2 * (EG + EH));
        bool exact = true;
        for (int dir = 0; dir < 2; ++dir) {
            uint16_t *i1 = b1.data() + dir * IMG1, *i2 = b2i.data() + dir * IMG2;
            exact &= tc_pack(w.W1.data() + (size_t)dir * EG * EH, EG, EH, EH, i1);
            exact &= tc_pack(w.R1.data() + (size_t)dir * EG * EH, EG, EH, 0, i1 + TcLayout<1>::BX_BYTES / 2);
            exact &= tc_pack(w.W2.data() + (size_t)dir * EG * 2 * EH, EG, 2 * EH, EH, i2);
            exact &= tc_pack(w.R2.data() + (size_t)dir * EG * EH, EG, EH, 0, i2 + TcLayout<2>::BX_BYTES / 2);
            for (int row = 0; row < EG; ++row) {
                bias1[(size_t)dir * (EG + EH) + row] = w.B1[(size_t)dir * 4 * EH + row];
                bias2[(size_t)dir * (EG + EH) + row] = b2[dir * EG + row];
            }
            for (int i = 0; i < EH; ++i) {
                bias1[(size_t)dir * (EG + EH) + EG + i] = bR1[dir * EH + i];
                bias2[(size_t)dir * (EG + EH) + EG + i] = bR2[dir * EH + i];
            }
        }
        // the 97 embedding rows the tokeniser can select (slot 0 = the pad id), as fp16: row-major, 128 bytes each
        std::vector<uint16_t> emb16((size_t)NSLOT * EH);
        for (int slot = 0; slot < NSLOT; ++slot) {
            const float *x = w.emb.data() + (size_t)(slot == 0 ? 0 : slot_token_id(slot - 1)) * EH;
            for (int k = 0; k < EH; ++k) {
                const __half hv = __float2half_rn(x[k]);
                if (__half2float(hv) != x[k]) exact = false;
                emb16[(size_t)slot * EH + k] = __half_as_ushort(hv);
            }
        }
        e.tc_exact = exact;
        if ((rc = upload(b1, &e.Bw1)) != DRM_OK) return rc;
        if ((rc = upload(b2i, &e.Bw2)) != DRM_OK) return rc;
        if ((rc = upload(bias1, &e.bias1)) != DRM_OK) return rc;
        if ((rc = upload(bias2, &e.bias2)) != DRM_OK) return rc;
        if ((rc = upload(emb16, &e.emb16)) != DRM_OK) return rc;
    }
#endif
    return DRM_OK;
}

void embedder_free(DeviceEmbedder &e)
{
    if (e.device >= 0) cudaSetDevice(e.device);
    for (float **p : {&e.P1, &e.R1, &e.bR1, &e.W2, &e.b2, &e.R2, &e.bR2, &e.Y1, &e.X2, &e.out, &e.bias1, &e.bias2}) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
    for (uint16_t **p : {&e.Bw1, &e.Bw2, &e.Y1h, &e.emb16}) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
    e.chunk = 0;
    e.chunk_tc = 0;
}

#ifndef DRM_EMU
// The tensor-core form: layer 1 (two CTAs per SM: 128 sequences x one direction each), then layer 2 with its input projection
// fused into the step's product (one CTA per SM).  Passes of sms x 128 x 2 sequences; scratch = the layer-1 states as fp16
// hi | lo pairs (63 KB per sequence, 2.4 GB per pass).
static int launch_embed_tc(DeviceEmbedder &e, const EmbedSource &src, int64_t first, int64_t n, float *out_dev, cudaStream_t stream, int sms)
{
    const int T = e.max_len;
    const int64_t full_pass = (int64_t)sms * TC_M * 2;
    const int64_t want = n < full_pass ? ((n + TC_M - 1) / TC_M) * TC_M : full_pass;
    if (want > e.chunk_tc) {
        DRM_CUDA_TRY(cudaStreamSynchronize(stream));
        if (e.Y1h) cudaFree(e.Y1h);
        e.Y1h = nullptr; e.chunk_tc = 0;
        DRM_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(&e.Y1h), (size_t)want * T * 4 * EH * sizeof(uint16_t)));
        e.chunk_tc = want;
    }
    SrcDev s;
    s.mode = src.mode; s.flat = src.flat; s.off = reinterpret_cast<const long long *>(src.off);
    s.cls = src.cls; s.form = src.form; s.ref_n = src.ref_n; s.ref_len = src.ref_len;
    s.first_id = src.first_id; s.id_step = src.id_step;
    s.ids = src.ids; s.decorate = src.decorate;
    // layer 1: 110 KB per CTA, two CTAs (2 x 256 TMEM columns) per SM; layer 2: 170 KB and all 512 columns, one CTA per SM
    constexpr int SM1 = TcLayout<1>::BYTES, SM2 = TcLayout<2>::BYTES;
    // threads per sequence in the gate phase, per layer: DRM_EMBED_TC_TPS = "<layer 1><layer 2>" (A/B; layer 1: 1 or 2, two CTAs per
    // SM; layer 2: 1, 2 or 4, one CTA per SM)
    int tps1 = 2, tps2 = 2;  // four threads per sequence in layer 2 measured no faster (the step is bound by its tensor products)
    if (const char *ev = getenv("DRM_EMBED_TC_TPS")) {
        if (ev[0] == '1') tps1 = 1;
        if (ev[0] && (ev[1] == '1' || ev[1] == '2' || ev[1] == '4')) tps2 = ev[1] - '0';
    }
    auto k1 = tps1 == 2 ? gru_tc_kernel<1, 2> : gru_tc_kernel<1, 1>;
    auto k2 = tps2 == 4 ? gru_tc_kernel<2, 4> : tps2 == 2 ? gru_tc_kernel<2, 2> : gru_tc_kernel<2, 1>;
    // DRM_EMBED_TC_XG (A/B): products of the next step's input half issued per group between the gates; default: all before them
    if (const char *xg = getenv("DRM_EMBED_TC_XG")) {
        if (tps2 == 2 && atoi(xg) == 2) k2 = gru_tc_kernel<2, 2, false, 2>;
        if (tps2 == 2 && atoi(xg) == 4) k2 = gru_tc_kernel<2, 2, false, 4>;
    }
    const char *pe = getenv("DRM_EMBED_TC_PHASES");
    const bool phases = pe && pe[0] == '1' && tps1 == 2 && tps2 == 2;
    if (phases) { k1 = gru_tc_kernel<1, 2, true>; k2 = gru_tc_kernel<2, 2, true>; }
    DRM_CUDA_TRY(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, SM1));
    DRM_CUDA_TRY(cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, SM2));
    // DRM_EMBED_TC_PHASES=1 (profiling only): CTA 0's clocks per phase of a step, printed after the (synchronised) call
    unsigned long long *phase_clk = nullptr;
    if (phases) {
        DRM_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(&phase_clk), 10 * sizeof(unsigned long long)));
        DRM_CUDA_TRY(cudaMemsetAsync(phase_clk, 0, 10 * sizeof(unsigned long long), stream));
    }
    for (int64_t done = 0; done < n; done += e.chunk_tc) {
        const int64_t m = (n - done) < e.chunk_tc ? (n - done) : e.chunk_tc;
        const unsigned blocks = (unsigned)((m + TC_M - 1) / TC_M) * 2;
        k1<<<blocks, TC_M * tps1, SM1, stream>>>(s, first + done, m, T, reinterpret_cast<const uint4 *>(e.emb16), reinterpret_cast<const uint4 *>(e.Bw1),
                                                  e.bias1, nullptr, reinterpret_cast<__half *>(e.Y1h), nullptr, phase_clk);
        DRM_CUDA_TRY(cudaGetLastError());
        k2<<<blocks, TC_M * tps2, SM2, stream>>>(s, first + done, m, T, nullptr, reinterpret_cast<const uint4 *>(e.Bw2), e.bias2,
                                                  reinterpret_cast<const __half *>(e.Y1h), nullptr, out_dev + (size_t)done * 2 * EH, phase_clk);
        DRM_CUDA_TRY(cudaGetLastError());
    }
    if (phase_clk) {
        unsigned long long h[10];
        DRM_CUDA_TRY(cudaStreamSynchronize(stream));
        DRM_CUDA_TRY(cudaMemcpy(h, phase_clk, sizeof(h), cudaMemcpyDeviceToHost));
        cudaFree(phase_clk);
        for (int l = 0; l < 2; ++l)
            fprintf(stderr, "[embed tc] layer %d, CTA 0, clocks per step: stage+barrier %.0f, issue %.0f, wait %.0f, restage %.0f, gates %.0f\n", l + 1,
                    (double)h[l * 5] / T, (double)h[l * 5 + 1] / T, (double)h[l * 5 + 2] / T, (double)h[l * 5 + 3] / T, (double)h[l * 5 + 4] / T);
    }
    return DRM_OK;
}
#endif

int launch_embed(DeviceEmbedder &e, const EmbedSource &src, int64_t first, int64_t n, float *out_dev, cudaStream_t stream)
{
    if (n <= 0) return DRM_OK;
    if (!e.P1) return fail(DRM_EINVAL, "embedder is not open");
    const int T = e.max_len;
    int sms = 148;
    DRM_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e.device));
#ifndef DRM_EMU
    {
        // default ON since it was measured (profiles/r2_embed_bench_tc.json: 5.1 M sequences/s against 1.05 M, outputs within 4e-6 of
        // the fp32 kernels); DRM_EMBED_TC=0 selects the fp32 kernels, which also serve weights that are not fp16 numbers
        const char *tc = getenv("DRM_EMBED_TC");
        const bool use_tc = e.tc_exact && !(tc && tc[0] == '0');
        if (use_tc) return launch_embed_tc(e, src, first, n, out_dev, stream, sms);
    }
#endif
    const int64_t full_pass = (int64_t)sms * ES * 4;  // 4 CTA waves
    const int64_t want = n < full_pass ? ((n + ES - 1) / ES) * ES : full_pass;
    if (want > e.chunk) {  // scratch grows to the largest pass seen (<= 2.4 GB)
        DRM_CUDA_TRY(cudaStreamSynchronize(stream));
        if (e.Y1) cudaFree(e.Y1);
        if (e.X2) cudaFree(e.X2);
        e.Y1 = e.X2 = nullptr; e.chunk = 0;
        DRM_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(&e.Y1), (size_t)want * T * 2 * EH * sizeof(float)));
        DRM_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(&e.X2), (size_t)want * T * 2 * EG * sizeof(float)));
        e.chunk = want;
    }
    SrcDev s;
    s.mode = src.mode; s.flat = src.flat; s.off = reinterpret_cast<const long long *>(src.off);
    s.cls = src.cls; s.form = src.form; s.ref_n = src.ref_n; s.ref_len = src.ref_len;
    s.first_id = src.first_id; s.id_step = src.id_step;
    s.ids = src.ids; s.decorate = src.decorate;
    // default ON: 1.04 M sequences/s against 0.72 M without (profiles/r2_embed_bench_prefetch.json); DRM_EMBED_PREFETCH=0 for A/B
    const char *pf = getenv("DRM_EMBED_PREFETCH");
    const bool prefetch = !(pf && pf[0] == '0');
    for (int64_t done = 0; done < n; done += e.chunk) {
        const int64_t m = (n - done) < e.chunk ? (n - done) : e.chunk;
        const unsigned blocks = (unsigned)((m + ES - 1) / ES);
        if (prefetch) gru_kernel<1, true><<<blocks, ETHREADS, 0, stream>>>(s, first + done, m, T, e.P1, nullptr, e.R1, e.bR1, e.Y1, nullptr);
        else gru_kernel<1, false><<<blocks, ETHREADS, 0, stream>>>(s, first + done, m, T, e.P1, nullptr, e.R1, e.bR1, e.Y1, nullptr);
        proj2_kernel<<<blocks, ETHREADS, 0, stream>>>(m, T, e.W2, e.b2, e.Y1, e.X2);
        float *dst = out_dev + (size_t)done * 2 * EH;
        if (prefetch) gru_kernel<2, true><<<blocks, ETHREADS, 0, stream>>>(s, first + done, m, T, nullptr, e.X2, e.R2, e.bR2, nullptr, dst);
        else gru_kernel<2, false><<<blocks, ETHREADS, 0, stream>>>(s, first + done, m, T, nullptr, e.X2, e.R2, e.bR2, nullptr, dst);
        DRM_CUDA_TRY(cudaGetLastError());
    }
    return DRM_OK;
}

}  // namespace drm
