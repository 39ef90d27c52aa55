// The embedder on the device: sequence bytes -> 3-mer tokens -> token embedding -> 2-layer
// bidirectional GRU (hidden 64) -> 128-d vector = [layer-2 forward final state | layer-2 backward final state].
// Replaces Vectorizer::vectorize (src/inference/vectorize.cpp:34-141: Preprocessor::preprocess src/inference/preprocess.cpp:20-42,
// prepareBatch :345-366, OpenVINO inference of models/finetuned_sgn33-new-a-Apr6.xml src/inference/fast_model.cpp:3-224).
// CPU restatement: oracle/embed_oracle.py.  fp32 arithmetic on the model's fp16-exact weights.
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

#include <cstdlib>

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
    return DRM_OK;
}

void embedder_free(DeviceEmbedder &e)
{
    if (e.device >= 0) cudaSetDevice(e.device);
    for (float **p : {&e.P1, &e.R1, &e.bR1, &e.W2, &e.b2, &e.R2, &e.bR2, &e.Y1, &e.X2, &e.out}) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
    e.chunk = 0;
}

int launch_embed(DeviceEmbedder &e, const EmbedSource &src, int64_t first, int64_t n, float *out_dev, cudaStream_t stream)
{
    if (n <= 0) return DRM_OK;
    if (!e.P1) return fail(DRM_EINVAL, "embedder is not open");
    const int T = e.max_len;
    int sms = 148;
    DRM_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e.device));
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
