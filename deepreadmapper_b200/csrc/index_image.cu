// FAISS arrays -> device image with FUSED rows: every adjacency row carries the PQ codes of its
// neighbours, so one hop of the search is one contiguous read instead of a row read followed by
// 2*M_hnsw dependent 8-byte gathers.  Costs (2*M_hnsw)*(4 + code_size) bytes per vertex (384 B at the
// reference's M_hnsw=16 / M_pq=8, i.e. 3.6 GB for the 9.28 M-vector E. coli index) - cheap on 180 GB.
// The gather itself runs on the GPU from the uploaded FAISS arrays.
#include "common.h"

#include <cstring>

namespace drm {

namespace {

__global__ void fill_rows0_kernel(const int32_t *__restrict__ neighbors, const unsigned long long *__restrict__ offsets,
                                  const uint8_t *__restrict__ codes, long long ntotal, int W0, int cs, int row_bytes,
                                  int code_off, uint8_t *__restrict__ rows0, int *__restrict__ err)
{
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long v = t / W0;
    int slot = (int)(t % W0);
    if (v >= ntotal) return;
    int id = neighbors[offsets[v] + slot];
    if (id >= ntotal) { atomicExch(err, 1); id = -1; }
    uint8_t *row = rows0 + (size_t)v * row_bytes;
    reinterpret_cast<int *>(row)[slot] = id < 0 ? -1 : id;
    uint8_t *dst = row + code_off + (size_t)slot * cs;
    if (id >= 0) {
        const uint8_t *src = codes + (size_t)id * cs;
        for (int b = 0; b < cs; ++b) dst[b] = src[b];
    } else {
        for (int b = 0; b < cs; ++b) dst[b] = 0;
    }
}

__global__ void fill_up_kernel(const int32_t *__restrict__ neighbors, const unsigned long long *__restrict__ offsets,
                               const uint8_t *__restrict__ codes, const int32_t *__restrict__ levels,
                               const int32_t *__restrict__ cum, const int32_t *__restrict__ up_base,
                               const int32_t *__restrict__ row_vertex, const int32_t *__restrict__ row_level,
                               long long n_rows, long long ntotal, int Wu, int cs, int row_bytes, int code_off,
                               uint8_t *__restrict__ up, int *__restrict__ err)
{
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long r = t / Wu;
    int slot = (int)(t % Wu);
    if (r >= n_rows) return;
    const int v = row_vertex[r], l = row_level[r];
    uint8_t *row = up + (size_t)r * row_bytes;
    if (slot == 0) {
        int *h = reinterpret_cast<int *>(row);
        h[0] = v; h[1] = l; h[2] = 0; h[3] = 0;
    }
    int width = cum[l + 1] - cum[l];
    int id = slot < width ? neighbors[offsets[v] + cum[l] + slot] : -1;
    int nrow = -1;
    if (id >= 0) {
        if (id >= ntotal || levels[id] <= l || up_base[id] < 0) { atomicExch(err, 2); id = -1; }
        else nrow = up_base[id] + (l - 1);
    }
    reinterpret_cast<int *>(row + 16)[slot] = nrow;
    uint8_t *dst = row + code_off + (size_t)slot * cs;
    if (id >= 0) {
        const uint8_t *src = codes + (size_t)id * cs;
        for (int b = 0; b < cs; ++b) dst[b] = src[b];
    } else {
        for (int b = 0; b < cs; ++b) dst[b] = 0;
    }
}

template <class T> int upload(T **dst, const void *src, size_t bytes)
{
    DRM_CUDA_TRY(cudaMalloc(dst, bytes ? bytes : 16));
    if (bytes) DRM_CUDA_TRY(cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice));
    return DRM_OK;
}

inline int round16(int x) { return (x + 15) / 16 * 16; }

}  // namespace

void free_device_index(DeviceIndex &dx)
{
    if (dx.device >= 0) cudaSetDevice(dx.device);
    cudaFree(dx.centroids); cudaFree(dx.centroids_t); cudaFree(dx.rows0); cudaFree(dx.up); cudaFree(dx.counters);
    dx.centroids = nullptr; dx.centroids_t = nullptr; dx.rows0 = nullptr; dx.up = nullptr; dx.counters = nullptr;
}

int build_device_index(const HostIndex &h, int device, DeviceIndex &dx)
{
    if (h.nbits != 8) return fail(DRM_EUNSUPPORTED, "only nbits == 8 IndexHNSWPQ files are supported on the GPU path");
    if (h.ntotal <= 0) return fail(DRM_EINVAL, "index is empty");
    if (h.ntotal >= ((int64_t)1 << 31) - 1) return fail(DRM_EUNSUPPORTED, "more than 2^31 vectors in one index shard");
    if (h.code_size > 64) return fail(DRM_EUNSUPPORTED, "PQ code size above 64 bytes");
    DRM_CUDA_TRY(cudaSetDevice(device));
    dx.device = device;
    dx.d = h.d; dx.M = (int)h.M; dx.ksub = (int)h.ksub; dx.dsub = (int)h.dsub; dx.code_size = (int)h.code_size;
    dx.ntotal = h.ntotal; dx.max_level = h.max_level;
    dx.W0 = h.cum[1]; dx.Wu = h.cum.size() > 2 ? h.cum[2] - h.cum[1] : h.cum[1] / 2;
    const int cs = dx.code_size;
    const int code0_off = round16(4 * dx.W0), codeu_off = round16(16 + 4 * dx.Wu);
    dx.code0_off = code0_off; dx.codeu_off = codeu_off;
    dx.row0_bytes = round16(code0_off + cs * dx.W0);
    dx.rowu_bytes = round16(codeu_off + cs * dx.Wu);

    // host side: row numbering of the upper levels
    std::vector<int32_t> up_base((size_t)h.ntotal, -1), row_vertex, row_level;
    int64_t n_rows = 0;
    for (int64_t v = 0; v < h.ntotal; ++v) {
        int lv = h.levels[(size_t)v];
        if (lv < 1 || (size_t)lv >= h.cum.size()) return fail(DRM_EIO, "index file: vertex level out of range");
        if (h.offsets[(size_t)v + 1] - h.offsets[(size_t)v] != (uint64_t)h.cum[(size_t)lv]) return fail(DRM_EIO, "index file: offsets do not match levels");
        if (lv > 1) {
            up_base[(size_t)v] = (int32_t)n_rows;
            for (int l = 1; l < lv; ++l) { row_vertex.push_back((int32_t)v); row_level.push_back(l); }
            n_rows += lv - 1;
        }
    }
    if (n_rows >= ((int64_t)1 << 31) - 1) return fail(DRM_EUNSUPPORTED, "too many upper-level rows");
    for (size_t l = 1; l + 1 < h.cum.size() && (int)l <= h.max_level; ++l)
        if (h.cum[l + 1] - h.cum[l] != dx.Wu) return fail(DRM_EUNSUPPORTED, "non-uniform upper-level row width");
    dx.n_up_rows = n_rows;
    dx.entry_vertex = h.entry_point;
    dx.entry_row = h.max_level >= 1 ? (int64_t)up_base[(size_t)h.entry_point] + h.max_level - 1 : -1;
    memcpy(dx.entry_code, &h.codes[(size_t)h.entry_point * cs], (size_t)cs);

    int32_t *d_nb = nullptr, *d_levels = nullptr, *d_cum = nullptr, *d_upb = nullptr, *d_rv = nullptr, *d_rl = nullptr;
    unsigned long long *d_off = nullptr;
    uint8_t *d_codes = nullptr;
    int *d_err = nullptr;
    int rc = DRM_OK;
    auto cleanup = [&]() {
        cudaFree(d_nb); cudaFree(d_levels); cudaFree(d_cum); cudaFree(d_upb); cudaFree(d_rv); cudaFree(d_rl);
        cudaFree(d_off); cudaFree(d_codes); cudaFree(d_err);
    };
#define STEP(x) do { rc = (x); if (rc != DRM_OK) { cleanup(); free_device_index(dx); return rc; } } while (0)
#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { cleanup(); free_device_index(dx); return fail(DRM_ECUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); } } while (0)
    STEP(upload(&dx.centroids, h.centroids.data(), h.centroids.size() * sizeof(float)));
    if (h.dsub % 4 == 0) {  // [m][j][t] -> [m][t/4][j][t%4]
        std::vector<float> ct(h.centroids.size());
        const size_t q4 = h.dsub / 4;
        for (size_t m = 0; m < h.M; ++m)
            for (size_t j = 0; j < h.ksub; ++j)
                for (size_t t = 0; t < h.dsub; ++t)
                    ct[((m * q4 + t / 4) * h.ksub + j) * 4 + t % 4] = h.centroids[(m * h.ksub + j) * h.dsub + t];
        STEP(upload(&dx.centroids_t, ct.data(), ct.size() * sizeof(float)));
    }
    STEP(upload(&d_nb, h.neighbors.data(), h.neighbors.size() * sizeof(int32_t)));
    STEP(upload(&d_off, h.offsets.data(), h.offsets.size() * sizeof(uint64_t)));
    STEP(upload(&d_codes, h.codes.data(), h.codes.size()));
    STEP(upload(&d_levels, h.levels.data(), h.levels.size() * sizeof(int32_t)));
    STEP(upload(&d_cum, h.cum.data(), h.cum.size() * sizeof(int32_t)));
    STEP(upload(&d_upb, up_base.data(), up_base.size() * sizeof(int32_t)));
    STEP(upload(&d_rv, row_vertex.data(), row_vertex.size() * sizeof(int32_t)));
    STEP(upload(&d_rl, row_level.data(), row_level.size() * sizeof(int32_t)));
    CU(cudaMalloc(&d_err, sizeof(int)));
    CU(cudaMemset(d_err, 0, sizeof(int)));
    CU(cudaMalloc(&dx.rows0, (size_t)h.ntotal * dx.row0_bytes));
    CU(cudaMalloc(&dx.up, (size_t)(n_rows ? n_rows : 1) * dx.rowu_bytes));
    CU(cudaMalloc(&dx.counters, 8 * sizeof(int64_t)));
    CU(cudaMemset(dx.rows0, 0, (size_t)h.ntotal * dx.row0_bytes));
    CU(cudaMemset(dx.up, 0, (size_t)(n_rows ? n_rows : 1) * dx.rowu_bytes));
    {
        long long threads = (long long)h.ntotal * dx.W0;
        fill_rows0_kernel<<<(unsigned)((threads + 255) / 256), 256>>>(d_nb, d_off, d_codes, h.ntotal, dx.W0, cs, dx.row0_bytes,
                                                                     code0_off, dx.rows0, d_err);
        CU(cudaGetLastError());
    }
    if (n_rows) {
        long long threads = n_rows * dx.Wu;
        fill_up_kernel<<<(unsigned)((threads + 255) / 256), 256>>>(d_nb, d_off, d_codes, d_levels, d_cum, d_upb, d_rv, d_rl, n_rows,
                                                                  h.ntotal, dx.Wu, cs, dx.rowu_bytes, codeu_off, dx.up, d_err);
        CU(cudaGetLastError());
    }
    CU(cudaDeviceSynchronize());
    int err = 0;
    CU(cudaMemcpy(&err, d_err, sizeof(int), cudaMemcpyDeviceToHost));
    cleanup();
#undef STEP
#undef CU
    if (err) { free_device_index(dx); return fail(DRM_EIO, "index file: neighbour id out of range or missing on its level"); }
    dx.bytes = h.centroids.size() * sizeof(float) + (size_t)h.ntotal * dx.row0_bytes + (size_t)n_rows * dx.rowu_bytes;
    return DRM_OK;
}

}  // namespace drm
