// Smith-Waterman with traceback: the alignment (reference interval, read interval, CIGAR operations, edit count)
// behind a rerank score.  SURVEY.md §8(f) rank 3: the reference's SAM writer carries a pseudo CIGAR "<len>M" and the
// note "TODO: Implement real CIGAR - from SW ranker" (src/utils/utils.cpp:381-384); this is that ranker-side piece.
//
// Recurrence and scoring = calc_sw_score (src/utils/metrics.cpp:10-45): +1 / -1 / -1, rows = window bases, columns =
// read bases, raw byte equality; window fetch and reverse complement = find_sequence (src/utils/post_processor.cpp:47-66).
// Conventions (restated on the CPU in oracle/sw_oracle.c: orc_sw_align):
//   * end cell = the first maximum in row-major order (the cell at which calc_sw_score's running maximum last rises);
//   * traceback while H > 0, preferring diagonal, then up (a window base without a read base: D), then left (I);
//   * operations are reported in alignment order as BAM-style words (length << 4 | op), op 0 = M, 1 = I, 2 = D;
//     soft clips are implied by the read interval.
// One warp per alignment: lane l owns read columns [l*CPL, (l+1)*CPL), window rows stream through the lanes as a
// 32-deep software pipeline (one shuffle per row), direction codes (2 bits per cell) go to shared memory, lane 0 walks
// them back.  22 500 cells per alignment against 2.88 M cells per read in the rerank: < 1 % of the stage.
#include "common.h"

namespace drm {

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int ALIGN_WARPS = 4;
constexpr int MAX_CPL = 8;  // read length <= 256

struct AlignDev {
    const uint8_t *raw; int form; long long ref_n; int ref_len;
    const char *q_flat; const long long *q_off;
    const unsigned long long *ids;
    long long n_align; int per_read, strip, max_ops;
    int *score, *ref_beg, *ref_end, *q_beg, *q_end, *n_ops, *nm;
    unsigned *ops;
};

__device__ __forceinline__ unsigned char comp_byte(unsigned char c)
{
    switch (c & 0x7f) {  // the reference's 128-entry complement table (src/utils/parse_inputs.cpp:5-14)
    case 'A': return 'T';
    case 'T': return 'A';
    case 'C': return 'G';
    case 'G': return 'C';
    case 'N': return 'N';
    default: return 0;
    }
}

__global__ void __launch_bounds__(ALIGN_WARPS * 32) sw_align_kernel(const AlignDev p)
{
    extern __shared__ unsigned short s_dir[];  // [warp][row][lane]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned short *dir = s_dir + (size_t)warp * p.ref_len * 32;
    const long long al = (long long)blockIdx.x * ALIGN_WARPS + warp;
    if (al >= p.n_align) return;
    const long long r = al / p.per_read;

    // window (find_sequence) and read (PREFIX / POSTFIX stripped like write_sam does)
    const unsigned long long id = p.ids[al];
    const uint8_t *a;
    int la = p.ref_len;
    bool rc = false;
    if (p.form == 1) {
        if (id >= (unsigned long long)p.ref_n) la = 0;
        a = p.raw + id * (unsigned long long)p.ref_len;
    } else {
        const unsigned long long pos = id >> 1;
        if (pos + (unsigned long long)p.ref_len > (unsigned long long)p.ref_n) la = 0;
        a = p.raw + pos;
        rc = (id & 1) != 0;
    }
    const uint8_t *b = reinterpret_cast<const uint8_t *>(p.q_flat) + p.q_off[r];
    int lb = (int)(p.q_off[r + 1] - p.q_off[r]);
    if (p.strip > 0 && lb > 2 * p.strip) { b += p.strip; lb -= 2 * p.strip; }
    const int CPL = (lb + 31) / 32;

    unsigned char bq[MAX_CPL];
    int H[MAX_CPL];
#pragma unroll
    for (int c = 0; c < MAX_CPL; ++c) {
        const int j = lane * CPL + c;
        bq[c] = (c < CPL && j < lb) ? b[j] : 0;
        H[c] = 0;
    }
    int best = 0, bi = 0, bj = 0, last_out = 0, prev_recv = 0;
    for (int t = 0; t < la + 31; ++t) {
        int recv = __shfl_up_sync(FULL, last_out, 1);
        if (lane == 0) recv = 0;
        const int row = t - lane;
        if (row >= 0 && row < la) {
            const unsigned char ca = rc ? comp_byte(a[la - 1 - row]) : a[row];
            int diag = prev_recv, left = recv;
            unsigned bits = 0;
#pragma unroll
            for (int c = 0; c < MAX_CPL; ++c) {
                const int j = lane * CPL + c;
                if (c < CPL && j < lb) {
                    const int up = H[c];
                    const int d = diag + (ca == bq[c] ? 1 : -1), u = up - 1, l = left - 1;
                    const int h = max(max(d, u), max(l, 0));
                    const unsigned code = h == 0 ? 0u : (h == d ? 1u : (h == u ? 2u : 3u));
                    bits |= code << (2 * c);
                    if (h > best) { best = h; bi = row; bj = j; }
                    diag = up;
                    left = h;
                    H[c] = h;
                }
            }
            last_out = left;
            dir[row * 32 + lane] = (unsigned short)bits;
        }
        prev_recv = recv;
    }
    __syncwarp();
    // first maximum in row-major order: largest score, then smallest row, then smallest column
    unsigned long long key = ((unsigned long long)(unsigned)best << 32) | ((unsigned long long)(0xffff - bi) << 16) | (unsigned)(0xffff - bj);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(FULL, key, o);
        key = other > key ? other : key;
    }
    if (lane != 0) return;
    best = (int)(key >> 32);
    int i = 0xffff - (int)((key >> 16) & 0xffff) + 1, j = 0xffff - (int)(key & 0xffff) + 1;  // 1-based end cell
    unsigned *ops = p.ops + (size_t)al * p.max_ops;
    int n_ops = 0, nm = 0, cur = -1, cur_len = 0;
    if (best == 0 || la == 0 || lb == 0) { i = 0; j = 0; }
    const int i_end = i, j_end = j;
    while (i > 0 && j > 0) {
        const int col = j - 1;
        const unsigned code = (dir[(i - 1) * 32 + col / CPL] >> (2 * (col % CPL))) & 3u;
        if (code == 0) break;
        const int op = code == 1 ? 0 : (code == 2 ? 2 : 1);
        if (code == 1) {
            const unsigned char ca = rc ? comp_byte(a[la - i]) : a[i - 1];
            if (ca != b[col]) ++nm;
            --i; --j;
        } else {
            ++nm;
            if (code == 2) --i; else --j;
        }
        if (op == cur) ++cur_len;
        else {
            if (cur >= 0) { if (n_ops < p.max_ops) ops[n_ops] = ((unsigned)cur_len << 4) | (unsigned)cur; ++n_ops; }
            cur = op; cur_len = 1;
        }
    }
    if (cur >= 0) { if (n_ops < p.max_ops) ops[n_ops] = ((unsigned)cur_len << 4) | (unsigned)cur; ++n_ops; }
    // the walk produced the operations end-to-start
    const int kept = n_ops < p.max_ops ? n_ops : p.max_ops;
    for (int x = 0, y = kept - 1; x < y; ++x, --y) { const unsigned tmp = ops[x]; ops[x] = ops[y]; ops[y] = tmp; }
    p.score[al] = best;
    p.ref_beg[al] = i; p.ref_end[al] = i_end;
    p.q_beg[al] = j; p.q_end[al] = j_end;
    p.n_ops[al] = n_ops;  // > max_ops: truncated (the caller can retry with a larger buffer)
    p.nm[al] = nm;
}

}  // namespace

// All pointers are device pointers.  ids: n_reads x per_read window ids; outputs: one entry per (read, slot), ops max_ops words each.
int launch_sw_align(const DeviceRef &ref, const unsigned long long *ids, int64_t n_reads, int per_read, const char *q_flat,
                    const int64_t *q_off, int max_qlen, int strip, int max_ops, int *score, int *ref_beg, int *ref_end, int *q_beg,
                    int *q_end, int *n_ops, int *nm, unsigned *ops, cudaStream_t stream)
{
    if (n_reads <= 0 || per_read <= 0) return DRM_OK;
    if (max_ops < 1) return fail(DRM_EINVAL, "max_ops must be >= 1");
    if (max_qlen - (max_qlen > 2 * strip ? 2 * strip : 0) > 32 * MAX_CPL) return fail(DRM_EUNSUPPORTED, "traceback handles reads up to 256 bases");
    if (ref.ref_len > 768) return fail(DRM_EUNSUPPORTED, "traceback handles windows up to 768 bases (direction codes live in shared memory)");
    AlignDev p;
    p.raw = ref.raw; p.form = ref.form; p.ref_n = ref.len; p.ref_len = ref.ref_len;
    p.q_flat = q_flat; p.q_off = reinterpret_cast<const long long *>(q_off);
    p.ids = ids; p.n_align = n_reads * per_read; p.per_read = per_read; p.strip = strip; p.max_ops = max_ops;
    p.score = score; p.ref_beg = ref_beg; p.ref_end = ref_end; p.q_beg = q_beg; p.q_end = q_end; p.n_ops = n_ops; p.nm = nm; p.ops = ops;
    const size_t smem = (size_t)ALIGN_WARPS * ref.ref_len * 32 * sizeof(unsigned short);
    DRM_CUDA_TRY(cudaFuncSetAttribute(sw_align_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long blocks = (p.n_align + ALIGN_WARPS - 1) / ALIGN_WARPS;
    sw_align_kernel<<<(unsigned)blocks, ALIGN_WARPS * 32, smem, stream>>>(p);
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

}  // namespace drm
