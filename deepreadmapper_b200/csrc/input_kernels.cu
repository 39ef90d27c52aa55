// Input side on the device: FASTA -> cleaned genome, FASTQ -> read sequences.
//
//   clean_fasta_device   extract_FASTA_sequence (src/utils/parse_inputs.cpp:174-220): drop the first line, keep only
//                        A/C/G/T/N after upper-casing (every other byte - white space, IUPAC codes, the '>' and the text
//                        of later header lines that is not one of those five letters - is dropped).
//   parse_fastq_device   format_fastq (src/utils/parse_inputs.cpp:843-950): lines are split at '\n' only (a '\r' stays in
//                        the sequence), line 4r + 1 of the file is read r, returned as '<' + line + '>' (PREFIX / POSTFIX,
//                        includes/utils/config.hpp); a last line without '\n' counts, a trailing '\n' opens no extra line.
//
// Both are stream compactions: a predicate count per 4 KB tile, an exclusive scan of the tile counts, and an ordered
// write - three coalesced passes over the file bytes, HBM-bound.
#include "common.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace drm {

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int TILE_THREADS = 256;
constexpr int BYTES_PER_THREAD = 16;
constexpr int TILE_BYTES = TILE_THREADS * BYTES_PER_THREAD;  // 4096

struct DevMem {  // device allocations of one call, freed together
    std::vector<void *> ptrs;
    ~DevMem() { for (void *q : ptrs) cudaFree(q); }
    template <class T> cudaError_t alloc(T **p, size_t bytes)
    {
        void *q = nullptr;
        cudaError_t e = cudaMalloc(&q, bytes ? bytes : 16);
        if (e == cudaSuccess) ptrs.push_back(q);
        *p = static_cast<T *>(q);
        return e;
    }
    void release(void *p) { ptrs.erase(std::remove(ptrs.begin(), ptrs.end(), p), ptrs.end()); }  // ownership moves out
};

// ------------------------------------------------------------------ exclusive scan of int64 (in place, hierarchical)
constexpr int SCAN_ITEMS = 4, SCAN_TILE = TILE_THREADS * SCAN_ITEMS;

// exclusive prefix of `v` over the block (TILE_THREADS threads); *total (nullable) receives the block sum
__device__ __forceinline__ long long block_exclusive_ll(long long v, long long *s_warp, long long *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    long long inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const long long t = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        long long w = lane < TILE_THREADS / 32 ? s_warp[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(FULL, w, o);
            if (lane >= o) w += t;
        }
        if (lane < TILE_THREADS / 32) s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const long long warp_base = warp ? s_warp[warp - 1] : 0;
    if (total) *total = s_warp[TILE_THREADS / 32 - 1];
    const long long r = warp_base + inc - v;
    __syncthreads();  // s_warp may be reused by the caller
    return r;
}

__global__ void __launch_bounds__(TILE_THREADS) scan_tile_kernel(long long *data, long long n, long long *tile_sums)
{
    __shared__ long long s_warp[TILE_THREADS / 32];
    const long long base = (long long)blockIdx.x * SCAN_TILE + (long long)threadIdx.x * SCAN_ITEMS;
    long long v[SCAN_ITEMS], sum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        v[i] = base + i < n ? data[base + i] : 0;
        sum += v[i];
    }
    long long total;
    long long run = block_exclusive_ll(sum, s_warp, &total);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        if (base + i < n) data[base + i] = run;
        run += v[i];
    }
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void scan_add_kernel(long long *data, long long n, const long long *tile_base)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) data[i] += tile_base[i / SCAN_TILE];
}

// data[i] <- sum of data[0 .. i)
int scan_exclusive(long long *data, long long n, cudaStream_t s, DevMem &mem)
{
    if (n <= 0) return DRM_OK;
    const long long nt = (n + SCAN_TILE - 1) / SCAN_TILE;
    long long *sums = nullptr;
    DRM_CUDA_TRY(mem.alloc(&sums, (size_t)nt * sizeof(long long)));
    scan_tile_kernel<<<(unsigned)nt, TILE_THREADS, 0, s>>>(data, n, sums);
    DRM_CUDA_TRY(cudaGetLastError());
    if (nt > 1) {
        int rc = scan_exclusive(sums, nt, s, mem);
        if (rc != DRM_OK) return rc;
        scan_add_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(data, n, sums);
        DRM_CUDA_TRY(cudaGetLastError());
    }
    return DRM_OK;
}

// ------------------------------------------------------------------ byte predicates
struct KeepGenome {  // extract_FASTA_sequence: after the first line, upper-cased A/C/G/T/N only
    long long seq_start;
    __device__ __forceinline__ bool operator()(unsigned char c, long long pos, unsigned char &out) const
    {
        unsigned char u = (c >= 'a' && c <= 'z') ? (unsigned char)(c - 32) : c;
        out = u;
        return pos >= seq_start && (u == 'A' || u == 'C' || u == 'G' || u == 'T' || u == 'N');
    }
};
struct KeepNewline {
    __device__ __forceinline__ bool operator()(unsigned char c, long long, unsigned char &out) const
    {
        out = c;
        return c == '\n';
    }
};

__global__ void first_newline_kernel(const unsigned char *__restrict__ data, long long size, unsigned long long *first)
{
    long long best = size;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < size; i += (long long)gridDim.x * blockDim.x)
        if (data[i] == '\n') { best = i; break; }  // positions visited by one thread ascend
    if (best < size) atomicMin(first, (unsigned long long)best);
}

template <class Pred>
__global__ void __launch_bounds__(TILE_THREADS) tile_count_kernel(const unsigned char *__restrict__ data, long long size, Pred pred,
                                                                  long long *__restrict__ tile_count)
{
    __shared__ long long s_warp[TILE_THREADS / 32];
    const long long p0 = (long long)blockIdx.x * TILE_BYTES + (long long)threadIdx.x * BYTES_PER_THREAD;
    int c = 0;
    for (int i = 0; i < BYTES_PER_THREAD; ++i) {
        unsigned char o;
        if (p0 + i < size && pred(data[p0 + i], p0 + i, o)) ++c;
    }
    long long total;
    block_exclusive_ll(c, s_warp, &total);
    if (threadIdx.x == 0) tile_count[blockIdx.x] = total;
}

// MODE 0: write the kept (transformed) bytes; MODE 1: write the positions of the kept bytes
template <class Pred, int MODE>
__global__ void __launch_bounds__(TILE_THREADS) tile_write_kernel(const unsigned char *__restrict__ data, long long size, Pred pred,
                                                                  const long long *__restrict__ tile_base,
                                                                  unsigned char *__restrict__ out_bytes, long long *__restrict__ out_pos)
{
    __shared__ long long s_warp[TILE_THREADS / 32];
    const long long p0 = (long long)blockIdx.x * TILE_BYTES + (long long)threadIdx.x * BYTES_PER_THREAD;
    unsigned char kept[BYTES_PER_THREAD];
    unsigned mask = 0;
    int c = 0;
    for (int i = 0; i < BYTES_PER_THREAD; ++i) {
        unsigned char o = 0;
        if (p0 + i < size && pred(data[p0 + i], p0 + i, o)) { mask |= 1u << i; ++c; }
        kept[i] = o;
    }
    long long w = tile_base[blockIdx.x] + block_exclusive_ll(c, s_warp, nullptr);
    for (int i = 0; i < BYTES_PER_THREAD; ++i)
        if (mask & (1u << i)) {
            if (MODE == 0) out_bytes[w] = kept[i];
            else out_pos[w] = p0 + i;
            ++w;
        }
}

// ------------------------------------------------------------------ FASTQ records
// Read r is line 4r + 1: bytes [nl[4r] + 1, 4r + 1 < n_nl ? nl[4r + 1] : size).  lens[r] = its length + 2 ('<', '>').
__global__ void fastq_len_kernel(const long long *__restrict__ nl, long long n_nl, long long size, long long n_reads,
                                 long long *__restrict__ lens, int *__restrict__ max_len)
{
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int len = 0;
    if (r < n_reads) {
        const long long k = 4 * r + 1;
        const long long b = nl[k - 1] + 1, e = k < n_nl ? nl[k] : size;
        len = (int)(e - b) + 2;
        lens[r] = len;
    } else if (r == n_reads) lens[r] = 0;  // the scan turns this slot into the total
    for (int o = 16; o > 0; o >>= 1) len = max(len, __shfl_xor_sync(FULL, len, o));
    if ((threadIdx.x & 31) == 0 && len > 0) atomicMax(max_len, len);
}

// one warp per read
__global__ void fastq_copy_kernel(const unsigned char *__restrict__ data, const long long *__restrict__ nl, long long n_reads,
                                  const long long *__restrict__ off, char *__restrict__ flat)
{
    const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const long long b = nl[4 * r] + 1, o0 = off[r];
    const int len = (int)(off[r + 1] - o0) - 2;
    if (lane == 0) { flat[o0] = '<'; flat[o0 + 1 + len] = '>'; }
    for (int i = lane; i < len; i += 32) flat[o0 + 1 + i] = (char)data[b + i];
}

// DRM_INPUT_VERBOSE=1: host->device copy and kernel times of one call on stderr (measurement aid).
struct PhaseTimer {
    bool on;
    cudaEvent_t e[3] = {nullptr, nullptr, nullptr};
    PhaseTimer() : on(getenv("DRM_INPUT_VERBOSE") != nullptr) { if (on) for (auto &x : e) cudaEventCreate(&x); }
    ~PhaseTimer() { if (on) for (auto &x : e) cudaEventDestroy(x); }
    void mark(int i) { if (on) cudaEventRecord(e[i], 0); }
    void report(const char *what, long long in_bytes, long long out_bytes)
    {
        if (!on) return;
        cudaEventSynchronize(e[2]);
        float h2d = 0, k = 0;
        cudaEventElapsedTime(&h2d, e[0], e[1]);
        cudaEventElapsedTime(&k, e[1], e[2]);
        fprintf(stderr, "[drm input] %s: %lld bytes in, %lld bytes out, H2D %.3f ms, kernels %.3f ms (%.1f GB/s of input)\n", what, in_bytes,
                out_bytes, h2d, k, k > 0 ? (double)in_bytes / k / 1e6 : 0.0);
    }
};

template <class Pred>
int compact(const unsigned char *d_data, long long size, Pred pred, int mode, unsigned char **out_bytes, long long **out_pos,
            long long *n_out, DevMem &mem)
{
    *n_out = 0;
    const long long nt = (size + TILE_BYTES - 1) / TILE_BYTES;
    if (nt == 0) return DRM_OK;
    long long *counts = nullptr;
    DRM_CUDA_TRY(mem.alloc(&counts, (size_t)(nt + 1) * sizeof(long long)));
    DRM_CUDA_TRY(cudaMemset(counts + nt, 0, sizeof(long long)));
    tile_count_kernel<Pred><<<(unsigned)nt, TILE_THREADS>>>(d_data, size, pred, counts);
    DRM_CUDA_TRY(cudaGetLastError());
    int rc = scan_exclusive(counts, nt + 1, 0, mem);
    if (rc != DRM_OK) return rc;
    long long total = 0;
    DRM_CUDA_TRY(cudaMemcpy(&total, counts + nt, sizeof(long long), cudaMemcpyDeviceToHost));
    *n_out = total;
    if (mode == 0) {
        DRM_CUDA_TRY(mem.alloc(out_bytes, (size_t)total));
        tile_write_kernel<Pred, 0><<<(unsigned)nt, TILE_THREADS>>>(d_data, size, pred, counts, *out_bytes, nullptr);
    } else {
        DRM_CUDA_TRY(mem.alloc(out_pos, (size_t)total * sizeof(long long)));
        tile_write_kernel<Pred, 1><<<(unsigned)nt, TILE_THREADS>>>(d_data, size, pred, counts, nullptr, *out_pos);
    }
    DRM_CUDA_TRY(cudaGetLastError());
    return DRM_OK;
}

}  // namespace

// FASTA file bytes (host) -> cleaned genome bytes in a fresh device buffer on `device` (*d_genome, caller frees with cudaFree).
int clean_fasta_device(const char *data, int64_t size, int device, uint8_t **d_genome, int64_t *len)
{
    *d_genome = nullptr; *len = 0;
    if (!data || size < 0) return fail(DRM_EINVAL, "bad FASTA buffer");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return fail(DRM_ECUDA, "no CUDA device: the device FASTA path has no CPU fallback"); }
    DRM_CUDA_TRY(cudaSetDevice(device));
    DevMem mem;
    unsigned char *d_data = nullptr;
    unsigned long long *d_first = nullptr;
    DRM_CUDA_TRY(mem.alloc(&d_data, (size_t)size));
    DRM_CUDA_TRY(mem.alloc(&d_first, sizeof(unsigned long long)));
    PhaseTimer pt;
    pt.mark(0);
    if (size) DRM_CUDA_TRY(cudaMemcpy(d_data, data, (size_t)size, cudaMemcpyHostToDevice));
    pt.mark(1);
    unsigned long long first = (unsigned long long)size;
    DRM_CUDA_TRY(cudaMemcpy(d_first, &first, sizeof first, cudaMemcpyHostToDevice));
    if (size) {
        first_newline_kernel<<<148 * 4, 256>>>(d_data, size, d_first);
        DRM_CUDA_TRY(cudaGetLastError());
    }
    DRM_CUDA_TRY(cudaMemcpy(&first, d_first, sizeof first, cudaMemcpyDeviceToHost));
    KeepGenome pred{first < (unsigned long long)size ? (long long)first + 1 : (long long)size};
    unsigned char *out = nullptr;
    long long n = 0;
    int rc = compact(d_data, size, pred, 0, &out, nullptr, &n, mem);
    if (rc != DRM_OK) return rc;
    pt.mark(2);
    DRM_CUDA_TRY(cudaDeviceSynchronize());
    pt.report("FASTA clean", size, n);
    if (!out) DRM_CUDA_TRY(mem.alloc(&out, 16));
    mem.release(out);
    *d_genome = out;
    *len = n;
    return DRM_OK;
}

// FASTQ file bytes (host) -> decorated read sequences on `device`: *d_flat / *d_off (n_reads + 1 offsets), caller frees both.
int parse_fastq_device(const char *data, int64_t size, int device, char **d_flat, int64_t **d_off, int64_t *n_reads,
                       int64_t *flat_bytes, int *max_len)
{
    *d_flat = nullptr; *d_off = nullptr; *n_reads = 0; *flat_bytes = 0; *max_len = 0;
    if (!data || size < 0) return fail(DRM_EINVAL, "bad FASTQ buffer");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return fail(DRM_ECUDA, "no CUDA device: the device FASTQ path has no CPU fallback"); }
    DRM_CUDA_TRY(cudaSetDevice(device));
    DevMem mem;
    unsigned char *d_data = nullptr;
    DRM_CUDA_TRY(mem.alloc(&d_data, (size_t)size));
    PhaseTimer pt;
    pt.mark(0);
    if (size) DRM_CUDA_TRY(cudaMemcpy(d_data, data, (size_t)size, cudaMemcpyHostToDevice));
    pt.mark(1);
    long long *nl = nullptr, n_nl = 0;
    int rc = compact(d_data, size, KeepNewline(), 1, nullptr, &nl, &n_nl, mem);
    if (rc != DRM_OK) return rc;
    const long long n_lines = n_nl + ((size > 0 && data[size - 1] != '\n') ? 1 : 0);
    const long long n = (n_lines + 2) / 4;
    long long *off = nullptr;
    int *d_max = nullptr;
    DRM_CUDA_TRY(mem.alloc(&off, (size_t)(n + 1) * sizeof(long long)));
    DRM_CUDA_TRY(mem.alloc(&d_max, sizeof(int)));
    DRM_CUDA_TRY(cudaMemset(d_max, 0, sizeof(int)));
    DRM_CUDA_TRY(cudaMemset(off, 0, (size_t)(n + 1) * sizeof(long long)));
    long long total = 0;
    if (n > 0) {
        fastq_len_kernel<<<(unsigned)((n + 1 + 255) / 256), 256>>>(nl, n_nl, size, n, off, d_max);
        DRM_CUDA_TRY(cudaGetLastError());
        rc = scan_exclusive(off, n + 1, 0, mem);
        if (rc != DRM_OK) return rc;
        DRM_CUDA_TRY(cudaMemcpy(&total, off + n, sizeof(long long), cudaMemcpyDeviceToHost));
    }
    char *flat = nullptr;
    DRM_CUDA_TRY(mem.alloc(&flat, (size_t)total));
    if (n > 0) {
        fastq_copy_kernel<<<(unsigned)((n * 32 + 255) / 256), 256>>>(d_data, nl, n, off, flat);
        DRM_CUDA_TRY(cudaGetLastError());
    }
    DRM_CUDA_TRY(cudaMemcpy(max_len, d_max, sizeof(int), cudaMemcpyDeviceToHost));
    pt.mark(2);
    DRM_CUDA_TRY(cudaDeviceSynchronize());
    pt.report("FASTQ reads", size, total);
    mem.release(flat); mem.release(off);
    *d_flat = flat; *d_off = reinterpret_cast<int64_t *>(off); *n_reads = n; *flat_bytes = total;
    return DRM_OK;
}

}  // namespace drm
