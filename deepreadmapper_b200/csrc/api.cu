// extern "C" boundary of libdrm_cuda.so (declared in include/drm_b200.h).  Host-pointer entry points
// copy in, run the kernels, copy out and report the three times separately; "_device" entry points
// only enqueue kernels on the caller's stream.
#include "common.h"

#include <algorithm>
#include <atomic>
#include <cstring>
#include <chrono>
#include <memory>
#include <mutex>
#include <thread>

namespace {

using namespace drm;

struct DevBuf {  // RAII device allocation on the current device
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    void release() { if (p) cudaFree(p); p = nullptr; }
    int alloc(size_t bytes) { DRM_CUDA_TRY(cudaMalloc(&p, bytes ? bytes : 16)); return DRM_OK; }
    template <class T> T *as() { return static_cast<T *>(p); }
};

int grow(void **p, size_t *cap, size_t bytes)
{
    if (*p && *cap >= bytes) return DRM_OK;
    if (*p) cudaFree(*p);
    *p = nullptr; *cap = 0;
    DRM_CUDA_TRY(cudaMalloc(p, bytes ? bytes : 16));
    *cap = bytes;
    return DRM_OK;
}

void free_workspace(StageWorkspace &w)
{
    for (auto &s : w.slot) {
        if (s.s) cudaStreamDestroy(s.s);
        cudaFree(s.q); cudaFree(s.D); cudaFree(s.I); cudaFree(s.qs); cudaFree(s.qo); cudaFree(s.sc); cudaFree(s.id);
        if (s.off) cudaFreeHost(s.off);
        s = StageSlot();
    }
    cudaFree(w.cnt); cudaFree(w.dmin);
    w.cnt = w.dmin = nullptr;
}

// Keep the stream-ordered allocator's memory across synchronisations (the rerank launcher takes its scratch from it).
void keep_pool_memory(int device)
{
    cudaMemPool_t mp;
    if (cudaDeviceGetDefaultMemPool(&mp, device) == cudaSuccess) {
        unsigned long long thr = ~0ull;
        cudaMemPoolSetAttribute(mp, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    cudaGetLastError();
}

struct Timer {
    cudaEvent_t a = nullptr, b = nullptr;
    Timer() { cudaEventCreate(&a); cudaEventCreate(&b); }
    ~Timer() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
    void start(cudaStream_t s) { cudaEventRecord(a, s); }
    void stop(cudaStream_t s) { cudaEventRecord(b, s); }
    double ms() { float t = 0; cudaEventSynchronize(b); cudaEventElapsedTime(&t, a, b); return t; }
};

// queries [q0, q1) of one host-pointer search call on one replica
int search_slice(const DeviceIndex &dx, const float *q, int64_t n, int d, int k, int ef, float *D, int64_t *I,
                 drm_search_stats *st)
{
    DRM_CUDA_TRY(cudaSetDevice(dx.device));
    cudaStream_t s;
    DRM_CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    DevBuf dq, dD, dI;
    int rc;
    if ((rc = dq.alloc((size_t)n * d * sizeof(float))) || (rc = dD.alloc((size_t)n * k * sizeof(float))) ||
        (rc = dI.alloc((size_t)n * k * sizeof(int64_t)))) { cudaStreamDestroy(s); return rc; }
    Timer th, tk, td;
    cudaMemsetAsync(dx.counters, 0, 3 * sizeof(int64_t), s);
    th.start(s);
    cudaMemcpyAsync(dq.p, q, (size_t)n * d * sizeof(float), cudaMemcpyHostToDevice, s);
    th.stop(s);
    tk.start(s);
    rc = launch_search(dx, dq.as<float>(), n, k, ef, dD.as<float>(), dI.as<int64_t>(), dx.counters, s);
    tk.stop(s);
    if (rc != DRM_OK) { cudaStreamDestroy(s); return rc; }
    td.start(s);
    cudaMemcpyAsync(D, dD.p, (size_t)n * k * sizeof(float), cudaMemcpyDeviceToHost, s);
    cudaMemcpyAsync(I, dI.p, (size_t)n * k * sizeof(int64_t), cudaMemcpyDeviceToHost, s);
    td.stop(s);
    int64_t cnt[3] = {0, 0, 0};
    cudaMemcpyAsync(cnt, dx.counters, sizeof cnt, cudaMemcpyDeviceToHost, s);
    cudaError_t e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { cudaStreamDestroy(s); return fail(DRM_ECUDA, std::string("search: ") + cudaGetErrorString(e)); }
    if (st) {
        st->nhops0 = cnt[0]; st->nhops_up = cnt[1]; st->ndis_eval = cnt[2];
        st->h2d_ms = th.ms(); st->device_ms = tk.ms(); st->d2h_ms = td.ms();
    }
    cudaStreamDestroy(s);
    return DRM_OK;
}

int open_ref(const char *bytes, int64_t len_bytes, int64_t len, int form, int ref_len, const int *devices, int n_dev, drm_ref **out)
{
    if (!out) return fail(DRM_EINVAL, "null output handle");
    *out = nullptr;
    if (!bytes || len < 0 || ref_len <= 0 || n_dev < 1 || !devices) return fail(DRM_EINVAL, "bad reference arguments");
    drm_ref *r = new drm_ref();
    static std::atomic<uint64_t> next_uid{1};
    r->uid = next_uid.fetch_add(1);
    for (int i = 0; i < n_dev; ++i) {
        DeviceRef dr;
        dr.device = devices[i]; dr.form = form; dr.ref_len = ref_len; dr.len = len;
        cudaError_t e = cudaSetDevice(dr.device);
        int *flag = nullptr;
        if (e == cudaSuccess) e = cudaMalloc(&dr.raw, len_bytes ? len_bytes : 16);
        if (e == cudaSuccess) e = cudaMalloc(&dr.data_alloc, (size_t)len_bytes + 2 * REF_PAD);
        if (e == cudaSuccess) e = cudaMemset(dr.data_alloc, 4, (size_t)len_bytes + 2 * REF_PAD);
        if (e == cudaSuccess) dr.data = dr.data_alloc + REF_PAD;
        if (e == cudaSuccess) e = cudaMalloc(&flag, sizeof(int));
        if (e == cudaSuccess) e = cudaMemset(flag, 0, sizeof(int));
        if (e == cudaSuccess && len_bytes) e = cudaMemcpy(dr.raw, bytes, (size_t)len_bytes, cudaMemcpyHostToDevice);
        int rc = DRM_OK;
        if (e == cudaSuccess) rc = launch_classify(dr.raw, dr.data, len_bytes, flag, 0);
        int other = 0;
        if (e == cudaSuccess && rc == DRM_OK) e = cudaMemcpy(&other, flag, sizeof(int), cudaMemcpyDeviceToHost);
        cudaFree(flag);
        dr.has_other = other != 0;
        r->rep.push_back(dr);
        if (e != cudaSuccess || rc != DRM_OK) {
            if (e != cudaSuccess) rc = fail(DRM_ECUDA, std::string("reference upload: ") + cudaGetErrorString(e));
            drm_ref_close(r);
            return rc;
        }
    }
    *out = r;
    return DRM_OK;
}

}  // namespace

static int search_rerank_impl(drm_index *ix, drm_ref *ref, int mode, const float *q, int64_t n, int d, int k_search, int ef,
                              const char *q_flat, const int64_t *q_off, const drm_reads *reads, int64_t stride, int64_t k,
                              int64_t k_clusters, float *D, int64_t *I, int32_t *sw_scores, uint64_t *sw_ids, drm_search_stats *stats,
                              double *timing_ms);

extern "C" {

const char *drm_last_error(void) { return drm::last_error_string().c_str(); }
const char *drm_version(void) { return "drm-b200 0.1 (sm_100a)"; }

int drm_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int drm_index_open(const char *index_file, const int *devices, int n_dev, drm_index **out)
{
    if (!out) return fail(DRM_EINVAL, "null output handle");
    *out = nullptr;
    if (!index_file || !devices || n_dev < 1) return fail(DRM_EINVAL, "bad arguments to drm_index_open");
    HostIndex h;
    int rc = read_index_file(index_file, h);
    if (rc != DRM_OK) return rc;
    if (drm_device_count() < 1) return fail(DRM_ECUDA, "no CUDA device: libdrm_cuda has no CPU fallback");
    drm_index *ix = new drm_index();
    for (int i = 0; i < n_dev; ++i) {
        DeviceIndex dx;
        rc = build_device_index(h, devices[i], dx);
        if (rc != DRM_OK) { drm_index_close(ix); return rc; }
        ix->rep.push_back(dx);
        ix->ws.emplace_back(new StageWorkspace());
        keep_pool_memory(devices[i]);
    }
    int64_t info[12] = {h.d, h.ntotal, (int64_t)h.M, (int64_t)h.nbits, h.M_hnsw(), h.entry_point, h.max_level,
                        h.efConstruction, h.efSearch, (int64_t)h.neighbors.size(), n_dev, (int64_t)ix->rep[0].bytes};
    memcpy(ix->info, info, sizeof info);
    *out = ix;
    return DRM_OK;
}

void drm_index_close(drm_index *ix)
{
    if (!ix) return;
    for (size_t i = 0; i < ix->rep.size(); ++i) {
        if (ix->rep[i].device >= 0) cudaSetDevice(ix->rep[i].device);
        if (i < ix->ws.size() && ix->ws[i]) free_workspace(*ix->ws[i]);
        free_device_index(ix->rep[i]);
    }
    delete ix;
}

int drm_index_info(const drm_index *ix, int64_t *info)
{
    if (!ix || !info) return fail(DRM_EINVAL, "null argument");
    memcpy(info, ix->info, sizeof ix->info);
    return DRM_OK;
}

int drm_search(drm_index *ix, const float *q, int64_t n, int d, int k, int ef, float *D, int64_t *I, drm_search_stats *stats)
{
    if (!ix) return fail(DRM_EINVAL, "null index");
    if (n <= 0 || !q) return fail(DRM_EEMPTY, "Query data is empty");
    if (d != ix->rep[0].d) return fail(DRM_EINVAL, "query dimension " + std::to_string(d) + " does not match index dimension " + std::to_string(ix->rep[0].d));
    if (!D || !I || k <= 0 || ef <= 0) return fail(DRM_EINVAL, "bad arguments to drm_search");
    const int R = (int)ix->rep.size();
    std::vector<drm_search_stats> st((size_t)R);
    std::vector<int> rcs((size_t)R, DRM_OK);
    std::vector<std::string> errs((size_t)R);
    auto work = [&](int g) {
        int64_t q0 = n * g / R, q1 = n * (g + 1) / R;
        memset(&st[(size_t)g], 0, sizeof(drm_search_stats));
        if (q1 <= q0) return;
        rcs[(size_t)g] = search_slice(ix->rep[(size_t)g], q + q0 * d, q1 - q0, d, k, ef, D + q0 * k, I + q0 * k, &st[(size_t)g]);
        if (rcs[(size_t)g] != DRM_OK) errs[(size_t)g] = drm::last_error_string();
    };
    if (R == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int g = 0; g < R; ++g) th.emplace_back(work, g);
        for (auto &t : th) t.join();
    }
    for (int g = 0; g < R; ++g)
        if (rcs[(size_t)g] != DRM_OK) return fail(rcs[(size_t)g], errs[(size_t)g]);
    if (stats) {
        memset(stats, 0, sizeof *stats);
        for (auto &s : st) {
            stats->nhops0 += s.nhops0; stats->nhops_up += s.nhops_up; stats->ndis_eval += s.ndis_eval;
            stats->device_ms = std::max(stats->device_ms, s.device_ms);
            stats->h2d_ms = std::max(stats->h2d_ms, s.h2d_ms);
            stats->d2h_ms = std::max(stats->d2h_ms, s.d2h_ms);
        }
    }
    return DRM_OK;
}

int drm_search_device(drm_index *ix, int slot, const float *q_dev, int64_t n, int d, int k, int ef, float *D_dev,
                      int64_t *I_dev, int64_t *counters_dev, void *stream)
{
    if (!ix || slot < 0 || slot >= (int)ix->rep.size()) return fail(DRM_EINVAL, "bad index handle or replica slot");
    if (d != ix->rep[(size_t)slot].d) return fail(DRM_EINVAL, "query dimension does not match index dimension");
    DRM_CUDA_TRY(cudaSetDevice(ix->rep[(size_t)slot].device));
    return launch_search(ix->rep[(size_t)slot], q_dev, n, k, ef, D_dev, I_dev, counters_dev, static_cast<cudaStream_t>(stream));
}

int drm_ref_open_genome(const char *genome, int64_t len, int ref_len, const int *devices, int n_dev, drm_ref **out)
{
    return open_ref(genome, len, len, 0, ref_len, devices, n_dev, out);
}

int drm_ref_open_windows(const char *windows_flat, int64_t n_windows, int ref_len, const int *devices, int n_dev, drm_ref **out)
{
    if (n_windows < 0 || ref_len <= 0) return fail(DRM_EINVAL, "bad window table");
    return open_ref(windows_flat, n_windows * ref_len, n_windows, 1, ref_len, devices, n_dev, out);
}

void drm_ref_close(drm_ref *r)
{
    if (!r) return;
    for (auto &dr : r->rep) {
        if (dr.device >= 0) cudaSetDevice(dr.device);
        cudaFree(dr.data_alloc); cudaFree(dr.raw);
    }
    delete r;
}

int drm_sw_pairs(const char *a_flat, const int64_t *a_off, const char *b_flat, const int64_t *b_off, int64_t n, int32_t *scores)
{
    if (n < 0 || (n > 0 && (!a_flat || !a_off || !b_flat || !b_off || !scores))) return fail(DRM_EINVAL, "bad arguments to drm_sw_pairs");
    if (n == 0) return DRM_OK;
    if (drm_device_count() < 1) return fail(DRM_ECUDA, "no CUDA device: libdrm_cuda has no CPU fallback");
    int64_t max_b = 0;
    for (int64_t i = 0; i < n; ++i) max_b = std::max(max_b, b_off[i + 1] - b_off[i]);
    const int64_t abytes = a_off[n], bbytes = b_off[n];
    DevBuf da, dao, db, dbo, ds;
    int rc;
    if ((rc = da.alloc((size_t)abytes)) || (rc = dao.alloc((size_t)(n + 1) * 8)) || (rc = db.alloc((size_t)bbytes)) ||
        (rc = dbo.alloc((size_t)(n + 1) * 8)) || (rc = ds.alloc((size_t)n * 4))) return rc;
    DRM_CUDA_TRY(cudaMemcpy(da.p, a_flat, (size_t)abytes, cudaMemcpyHostToDevice));
    DRM_CUDA_TRY(cudaMemcpy(dao.p, a_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice));
    DRM_CUDA_TRY(cudaMemcpy(db.p, b_flat, (size_t)bbytes, cudaMemcpyHostToDevice));
    DRM_CUDA_TRY(cudaMemcpy(dbo.p, b_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice));
    rc = launch_sw_pairs(da.as<char>(), dao.as<int64_t>(), db.as<char>(), dbo.as<int64_t>(), n, (int)max_b, ds.as<int32_t>(), 0);
    if (rc != DRM_OK) return rc;
    DRM_CUDA_TRY(cudaMemcpy(scores, ds.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    return DRM_OK;
}

int drm_sw_rerank_device(drm_ref *r, int slot, int mode, const int64_t *neighbors_dev, int64_t n, int64_t k_in,
                         const char *q_flat_dev, const int64_t *q_off_dev, int max_qlen, int64_t stride, int64_t k,
                         int64_t k_clusters, int32_t *out_scores_dev, uint64_t *out_ids_dev, int32_t *n_cands_dev,
                         int32_t *min_cands_dev, int64_t *cells_dev, void *stream)
{
    if (!r || slot < 0 || slot >= (int)r->rep.size()) return fail(DRM_EINVAL, "bad reference handle or replica slot");
    const DeviceRef &dr = r->rep[(size_t)slot];
    DRM_CUDA_TRY(cudaSetDevice(dr.device));
    RerankArgs a{mode, neighbors_dev, n, k_in, q_flat_dev, q_off_dev, max_qlen, stride, k, k_clusters,
                 out_scores_dev, out_ids_dev, n_cands_dev, min_cands_dev, cells_dev};
    return launch_rerank(dr, a, static_cast<cudaStream_t>(stream));
}

int drm_sw_rerank(drm_ref *r, int mode, const int64_t *neighbors, int64_t n, int64_t k_in, const char *q_flat,
                  const int64_t *q_off, int64_t stride, int64_t k, int64_t k_clusters, int32_t *out_scores,
                  uint64_t *out_ids, int32_t *n_cands, double *device_ms)
{
    if (!r) return fail(DRM_EINVAL, "null reference handle");
    if (n < 0 || (n > 0 && (!neighbors || !q_flat || !q_off || !out_scores || !out_ids))) return fail(DRM_EINVAL, "bad arguments to drm_sw_rerank");
    if (device_ms) *device_ms = 0;
    if (n == 0) return DRM_OK;
    const int R = (int)r->rep.size();
    std::vector<int> rcs((size_t)R, DRM_OK);
    std::vector<std::string> errs((size_t)R);
    std::vector<double> ms((size_t)R, 0.0);
    std::vector<int> mins((size_t)R, 0x7fffffff);
    auto work = [&](int g) {
        int64_t r0 = n * g / R, r1 = n * (g + 1) / R, m = r1 - r0;
        if (m <= 0) return;
        auto body = [&]() -> int {
            const DeviceRef &dr = r->rep[(size_t)g];
            DRM_CUDA_TRY(cudaSetDevice(dr.device));
            int64_t qb = q_off[r0], qe = q_off[r1];
            int max_qlen = 0;
            std::vector<int64_t> off((size_t)m + 1);
            for (int64_t i = 0; i <= m; ++i) off[(size_t)i] = q_off[r0 + i] - qb;
            for (int64_t i = 0; i < m; ++i) max_qlen = std::max<int64_t>(max_qlen, off[(size_t)i + 1] - off[(size_t)i]);
            DevBuf dn, dq, dqo, dsc, did, dnc, dmin;
            int rc;
            if ((rc = dn.alloc((size_t)m * k_in * 8)) || (rc = dq.alloc((size_t)(qe - qb))) || (rc = dqo.alloc((size_t)(m + 1) * 8)) ||
                (rc = dsc.alloc((size_t)m * k * 4)) || (rc = did.alloc((size_t)m * k * 8)) || (rc = dnc.alloc((size_t)m * 4)) ||
                (rc = dmin.alloc(4))) return rc;
            cudaStream_t s;
            DRM_CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
            int big = 0x7fffffff;
            cudaMemcpyAsync(dn.p, neighbors + r0 * k_in, (size_t)m * k_in * 8, cudaMemcpyHostToDevice, s);
            cudaMemcpyAsync(dq.p, q_flat + qb, (size_t)(qe - qb), cudaMemcpyHostToDevice, s);
            cudaMemcpyAsync(dqo.p, off.data(), (size_t)(m + 1) * 8, cudaMemcpyHostToDevice, s);
            cudaMemcpyAsync(dmin.p, &big, 4, cudaMemcpyHostToDevice, s);
            Timer t;
            t.start(s);
            RerankArgs a{mode, dn.as<int64_t>(), m, k_in, dq.as<char>(), dqo.as<int64_t>(), max_qlen, stride, k, k_clusters,
                         dsc.as<int32_t>(), did.as<uint64_t>(), dnc.as<int32_t>(), dmin.as<int32_t>(), nullptr};
            rc = launch_rerank(dr, a, s);
            t.stop(s);
            if (rc == DRM_OK) {
                cudaMemcpyAsync(out_scores + r0 * k, dsc.p, (size_t)m * k * 4, cudaMemcpyDeviceToHost, s);
                cudaMemcpyAsync(out_ids + r0 * k, did.p, (size_t)m * k * 8, cudaMemcpyDeviceToHost, s);
                if (n_cands) cudaMemcpyAsync(n_cands + r0, dnc.p, (size_t)m * 4, cudaMemcpyDeviceToHost, s);
                cudaMemcpyAsync(&mins[(size_t)g], dmin.p, 4, cudaMemcpyDeviceToHost, s);
            }
            cudaError_t e = cudaStreamSynchronize(s);
            if (rc == DRM_OK) ms[(size_t)g] = t.ms();
            cudaStreamDestroy(s);
            if (rc != DRM_OK) return rc;
            if (e != cudaSuccess) return fail(DRM_ECUDA, std::string("rerank: ") + cudaGetErrorString(e));
            return DRM_OK;
        };
        rcs[(size_t)g] = body();
        if (rcs[(size_t)g] != DRM_OK) errs[(size_t)g] = drm::last_error_string();
    };
    if (R == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int g = 0; g < R; ++g) th.emplace_back(work, g);
        for (auto &t : th) t.join();
    }
    int min_c = 0x7fffffff;
    for (int g = 0; g < R; ++g) {
        if (rcs[(size_t)g] != DRM_OK) return fail(rcs[(size_t)g], errs[(size_t)g]);
        min_c = std::min(min_c, mins[(size_t)g]);
        if (device_ms) *device_ms = std::max(*device_ms, ms[(size_t)g]);
    }
    if (k > 0 && min_c < k)
        return fail(DRM_ENOTENOUGH, "Not enough candidates (" + std::to_string(min_c) + " < " + std::to_string(k) + ")");
    return DRM_OK;
}

int drm_search_rerank(drm_index *ix, drm_ref *ref, int mode, const float *q, int64_t n, int d, int k_search, int ef,
                      const char *q_flat, const int64_t *q_off, int64_t stride, int64_t k, int64_t k_clusters, float *D,
                      int64_t *I, int32_t *sw_scores, uint64_t *sw_ids, drm_search_stats *stats, double *timing_ms)
{
    if (!q_flat || !q_off) return fail(DRM_EINVAL, "query sequences are required for the SW rerank");
    return search_rerank_impl(ix, ref, mode, q, n, d, k_search, ef, q_flat, q_off, nullptr, stride, k, k_clusters, D, I, sw_scores,
                              sw_ids, stats, timing_ms);
}

int drm_search_rerank_reads(drm_index *ix, drm_ref *ref, int mode, const float *q, int64_t n, int d, int k_search, int ef,
                            const drm_reads *reads, int64_t stride, int64_t k, int64_t k_clusters, float *D, int64_t *I,
                            int32_t *sw_scores, uint64_t *sw_ids, drm_search_stats *stats, double *timing_ms)
{
    if (!reads) return fail(DRM_EINVAL, "null reads handle");
    if (n > reads->n) return fail(DRM_EINVAL, "more query embeddings than parsed reads");
    return search_rerank_impl(ix, ref, mode, q, n, d, k_search, ef, nullptr, nullptr, reads, stride, k, k_clusters, D, I, sw_scores,
                              sw_ids, stats, timing_ms);
}

}  // extern "C"

// One implementation for both forms of the read sequences: host bytes + offsets (uploaded chunk by chunk), or a
// drm_reads handle (already on every replica's device: nothing but the embeddings goes up).
static int search_rerank_impl(drm_index *ix, drm_ref *ref, int mode, const float *q, int64_t n, int d, int k_search, int ef,
                              const char *q_flat, const int64_t *q_off, const drm_reads *reads, int64_t stride, int64_t k,
                              int64_t k_clusters, float *D, int64_t *I, int32_t *sw_scores, uint64_t *sw_ids, drm_search_stats *stats,
                              double *timing_ms)
{
    if (!ix || !ref) return fail(DRM_EINVAL, "null handle");
    if (n <= 0 || !q) return fail(DRM_EEMPTY, "Query data is empty");
    if (reads && reads->rep.size() != ix->rep.size()) return fail(DRM_EINVAL, "reads and index must be opened on the same devices");
    if (d != ix->rep[0].d) return fail(DRM_EINVAL, "query dimension does not match index dimension");
    if (ix->rep.size() != ref->rep.size()) return fail(DRM_EINVAL, "index and reference must be opened on the same devices");
    const int R = (int)ix->rep.size();
    // Reads per pipeline chunk.  Small enough that the first upload / last download (not overlapped) are a few per cent
    // of the call, large enough that the persistent search grid (3552 warps) sees >= 9 queries per warp.
    int64_t CH = 32768;
    if (const char *e = getenv("DRM_CHUNK")) { long long v = atoll(e); if (v >= 1024 && v <= (1 << 22)) CH = v; }
    std::vector<int> rcs((size_t)R, DRM_OK), mins((size_t)R, 0x7fffffff);
    std::vector<std::string> errs((size_t)R);
    std::vector<drm_search_stats> st((size_t)R);
    std::vector<double> span((size_t)R, 0.0);
    auto t_begin = std::chrono::steady_clock::now();
    auto work = [&](int g) {
        const int64_t r0 = n * g / R, r1 = n * (g + 1) / R;
        memset(&st[(size_t)g], 0, sizeof(drm_search_stats));
        if (r1 <= r0) return;
        auto body = [&]() -> int {
            const DeviceIndex &dx = ix->rep[(size_t)g];
            const DeviceRef &dr = ref->rep[(size_t)g];
            if (dx.device != dr.device) return fail(DRM_EINVAL, "index and reference replicas are on different devices");
            DRM_CUDA_TRY(cudaSetDevice(dx.device));
            StageWorkspace &w = *ix->ws[(size_t)g];
            std::lock_guard<std::mutex> lock(w.mu);
            int64_t max_qbytes = 0;
            int max_qlen = 0;
            if (reads) {
                if (reads->rep[(size_t)g].device != dx.device) return fail(DRM_EINVAL, "reads and index replicas are on different devices");
                max_qlen = reads->max_len;
            } else {
                for (int64_t c0 = r0; c0 < r1; c0 += CH) {
                    int64_t c1 = std::min(r1, c0 + CH);
                    max_qbytes = std::max(max_qbytes, q_off[c1] - q_off[c0]);
                }
                for (int64_t i = r0; i < r1; ++i) max_qlen = std::max<int64_t>(max_qlen, q_off[i + 1] - q_off[i]);
            }
            const int64_t chn = std::min<int64_t>(CH, r1 - r0);
            int rc;
            if (!w.cnt) DRM_CUDA_TRY(cudaMalloc(&w.cnt, 4 * sizeof(int64_t)));
            if (!w.dmin) DRM_CUDA_TRY(cudaMalloc(&w.dmin, 16));
            const bool regrow_off = w.cap_qo < (size_t)(chn + 1) * 8;
            for (auto &s : w.slot) {
                if (!s.s) DRM_CUDA_TRY(cudaStreamCreateWithFlags(&s.s, cudaStreamNonBlocking));
                size_t cq = w.cap_q, cD = w.cap_D, cI = w.cap_I, cqs = w.cap_qs, cqo = w.cap_qo, csc = w.cap_sc, cid = w.cap_id;
                if ((rc = grow(&s.q, &cq, (size_t)chn * d * 4)) || (rc = grow(&s.D, &cD, (size_t)chn * k_search * 4)) ||
                    (rc = grow(&s.I, &cI, (size_t)chn * k_search * 8)) || (rc = grow(&s.qs, &cqs, (size_t)max_qbytes)) ||
                    (rc = grow(&s.qo, &cqo, (size_t)(chn + 1) * 8)) || (rc = grow(&s.sc, &csc, (size_t)chn * k * 4)) ||
                    (rc = grow(&s.id, &cid, (size_t)chn * k * 8))) return rc;
                if (regrow_off || !s.off) {
                    if (s.off) cudaFreeHost(s.off);
                    s.off = nullptr;
                    DRM_CUDA_TRY(cudaHostAlloc(&s.off, (size_t)(chn + 1) * 8, cudaHostAllocDefault));
                }
                if (&s == &w.slot[StageWorkspace::NSLOT - 1]) {
                    w.cap_q = cq; w.cap_D = cD; w.cap_I = cI; w.cap_qs = cqs; w.cap_qo = cqo; w.cap_sc = csc; w.cap_id = cid;
                }
            }
            cudaStream_t s0 = w.slot[0].s;
            int big = 0x7fffffff;
            DRM_CUDA_TRY(cudaMemsetAsync(w.cnt, 0, 4 * sizeof(int64_t), s0));
            DRM_CUDA_TRY(cudaMemcpyAsync(w.dmin, &big, 4, cudaMemcpyHostToDevice, s0));
            DRM_CUDA_TRY(cudaStreamSynchronize(s0));
            Timer span_t;
            span_t.start(s0);
            int which = 0;
            for (int64_t c0 = r0; c0 < r1; c0 += CH, which = (which + 1) % StageWorkspace::NSLOT) {
                StageSlot &s = w.slot[which];
                const int64_t m = std::min(r1, c0 + CH) - c0;
                DRM_CUDA_TRY(cudaStreamSynchronize(s.s));  // the slot's previous chunk (and its host offsets) are done
                cudaMemcpyAsync(s.q, q + c0 * d, (size_t)m * d * 4, cudaMemcpyHostToDevice, s.s);
                const char *seq_dev = static_cast<char *>(s.qs);
                const int64_t *off_dev = static_cast<int64_t *>(s.qo);
                if (reads) {  // absolute offsets into the resident buffer
                    seq_dev = reads->rep[(size_t)g].flat;
                    off_dev = reads->rep[(size_t)g].off + c0;
                } else {
                    const int64_t qb = q_off[c0];
                    for (int64_t i = 0; i <= m; ++i) s.off[i] = q_off[c0 + i] - qb;
                    cudaMemcpyAsync(s.qs, q_flat + qb, (size_t)(q_off[c0 + m] - qb), cudaMemcpyHostToDevice, s.s);
                    cudaMemcpyAsync(s.qo, s.off, (size_t)(m + 1) * 8, cudaMemcpyHostToDevice, s.s);
                }
                rc = launch_search(dx, static_cast<float *>(s.q), m, k_search, ef, static_cast<float *>(s.D), static_cast<int64_t *>(s.I),
                                   static_cast<int64_t *>(w.cnt), s.s);
                if (rc != DRM_OK) return rc;
                RerankArgs a{mode, static_cast<int64_t *>(s.I), m, k_search, seq_dev, off_dev, max_qlen,
                             stride, k, k_clusters, static_cast<int32_t *>(s.sc), static_cast<uint64_t *>(s.id), nullptr,
                             static_cast<int32_t *>(w.dmin), nullptr};
                rc = launch_rerank(dr, a, s.s);
                if (rc != DRM_OK) return rc;
                if (D) cudaMemcpyAsync(D + c0 * k_search, s.D, (size_t)m * k_search * 4, cudaMemcpyDeviceToHost, s.s);
                if (I) cudaMemcpyAsync(I + c0 * k_search, s.I, (size_t)m * k_search * 8, cudaMemcpyDeviceToHost, s.s);
                if (sw_scores) cudaMemcpyAsync(sw_scores + c0 * k, s.sc, (size_t)m * k * 4, cudaMemcpyDeviceToHost, s.s);
                if (sw_ids) cudaMemcpyAsync(sw_ids + c0 * k, s.id, (size_t)m * k * 8, cudaMemcpyDeviceToHost, s.s);
            }
            cudaError_t e = cudaSuccess;
            for (int i = StageWorkspace::NSLOT - 1; i >= 0; --i) {
                cudaError_t ei = cudaStreamSynchronize(w.slot[i].s);
                if (e == cudaSuccess) e = ei;
            }
            span_t.stop(s0);
            if (e != cudaSuccess) return fail(DRM_ECUDA, std::string("search_rerank: ") + cudaGetErrorString(e));
            span[(size_t)g] = span_t.ms();
            int64_t c[4];
            DRM_CUDA_TRY(cudaMemcpy(c, w.cnt, sizeof c, cudaMemcpyDeviceToHost));
            DRM_CUDA_TRY(cudaMemcpy(&mins[(size_t)g], w.dmin, 4, cudaMemcpyDeviceToHost));
            st[(size_t)g].nhops0 = c[0]; st[(size_t)g].nhops_up = c[1]; st[(size_t)g].ndis_eval = c[2];
            return DRM_OK;
        };
        rcs[(size_t)g] = body();
        if (rcs[(size_t)g] != DRM_OK) errs[(size_t)g] = drm::last_error_string();
    };
    if (R == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int g = 0; g < R; ++g) th.emplace_back(work, g);
        for (auto &t : th) t.join();
    }
    int min_c = 0x7fffffff;
    if (stats) memset(stats, 0, sizeof *stats);
    double dev_span = 0;
    for (int g = 0; g < R; ++g) {
        if (rcs[(size_t)g] != DRM_OK) return fail(rcs[(size_t)g], errs[(size_t)g]);
        min_c = std::min(min_c, mins[(size_t)g]);
        dev_span = std::max(dev_span, span[(size_t)g]);
        if (stats) { stats->nhops0 += st[(size_t)g].nhops0; stats->nhops_up += st[(size_t)g].nhops_up; stats->ndis_eval += st[(size_t)g].ndis_eval; }
    }
    if (stats) stats->device_ms = dev_span;
    if (timing_ms) {
        timing_ms[0] = dev_span;
        timing_ms[1] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
    }
    if (k > 0 && min_c < k)
        return fail(DRM_ENOTENOUGH, "Not enough candidates (" + std::to_string(min_c) + " < " + std::to_string(k) + ")");
    return DRM_OK;
}

extern "C" {

/* ------------------------------------------------------------------ SW traceback */
int drm_sw_align(drm_ref *r, const uint64_t *ids, int64_t n, int per_read, const char *q_flat, const int64_t *q_off, int strip,
                 int max_ops, int32_t *score, int32_t *ref_beg, int32_t *ref_end, int32_t *q_beg, int32_t *q_end, int32_t *n_ops,
                 int32_t *nm, uint32_t *ops)
{
    if (!r || r->rep.empty()) return fail(DRM_EINVAL, "null reference handle");
    if (n <= 0) return DRM_OK;
    if (!ids || !q_flat || !q_off || per_read < 1 || max_ops < 1 || !score || !ref_beg || !ref_end || !q_beg || !q_end || !n_ops || !nm || !ops)
        return fail(DRM_EINVAL, "bad arguments to drm_sw_align");
    const DeviceRef &dr = r->rep[0];
    DRM_CUDA_TRY(cudaSetDevice(dr.device));
    const int64_t na = n * per_read, qbytes = q_off[n] - q_off[0];
    int max_qlen = 0;
    for (int64_t i = 0; i < n; ++i) max_qlen = std::max<int64_t>(max_qlen, q_off[i + 1] - q_off[i]);
    std::vector<int64_t> rel((size_t)n + 1);
    for (int64_t i = 0; i <= n; ++i) rel[(size_t)i] = q_off[i] - q_off[0];
    DevBuf d_ids, d_q, d_off, d_out, d_ops;
    int rc;
    if ((rc = d_ids.alloc((size_t)na * 8)) || (rc = d_q.alloc((size_t)qbytes)) || (rc = d_off.alloc((size_t)(n + 1) * 8)) ||
        (rc = d_out.alloc((size_t)na * 7 * sizeof(int))) || (rc = d_ops.alloc((size_t)na * max_ops * sizeof(unsigned)))) return rc;
    DRM_CUDA_TRY(cudaMemcpy(d_ids.p, ids, (size_t)na * 8, cudaMemcpyHostToDevice));
    if (qbytes) DRM_CUDA_TRY(cudaMemcpy(d_q.p, q_flat + q_off[0], (size_t)qbytes, cudaMemcpyHostToDevice));
    DRM_CUDA_TRY(cudaMemcpy(d_off.p, rel.data(), (size_t)(n + 1) * 8, cudaMemcpyHostToDevice));
    DRM_CUDA_TRY(cudaMemset(d_ops.p, 0, (size_t)na * max_ops * sizeof(unsigned)));
    int *o = d_out.as<int>();
    rc = launch_sw_align(dr, d_ids.as<unsigned long long>(), n, per_read, d_q.as<char>(), d_off.as<int64_t>(), max_qlen, strip, max_ops, o,
                         o + na, o + 2 * na, o + 3 * na, o + 4 * na, o + 5 * na, o + 6 * na, d_ops.as<unsigned>(), 0);
    if (rc != DRM_OK) return rc;
    int32_t *dst[7] = {score, ref_beg, ref_end, q_beg, q_end, n_ops, nm};
    for (int f = 0; f < 7; ++f) DRM_CUDA_TRY(cudaMemcpy(dst[f], o + (size_t)f * na, (size_t)na * sizeof(int), cudaMemcpyDeviceToHost));
    DRM_CUDA_TRY(cudaMemcpy(ops, d_ops.p, (size_t)na * max_ops * sizeof(unsigned), cudaMemcpyDeviceToHost));
    return DRM_OK;
}

/* ------------------------------------------------------------------ input side on the device */
int drm_fasta_clean(const char *data, int64_t size, int device, char *out, int64_t cap, int64_t *len)
{
    if (!len) return fail(DRM_EINVAL, "null length output");
    uint8_t *g = nullptr;
    int rc = clean_fasta_device(data, size, device, &g, len);
    if (rc != DRM_OK) return rc;
    cudaError_t e = cudaSuccess;
    if (out && *len <= cap && *len > 0) e = cudaMemcpy(out, g, (size_t)*len, cudaMemcpyDeviceToHost);
    cudaFree(g);
    if (e != cudaSuccess) return fail(DRM_ECUDA, std::string("drm_fasta_clean: ") + cudaGetErrorString(e));
    if (out && *len > cap) return fail(DRM_EINVAL, "output buffer too small for the cleaned genome");
    return DRM_OK;
}

int drm_ref_open_fasta(const char *data, int64_t size, int ref_len, const int *devices, int n_dev, drm_ref **out, int64_t *genome_len)
{
    if (!out) return fail(DRM_EINVAL, "null output handle");
    *out = nullptr;
    if (!devices || n_dev < 1) return fail(DRM_EINVAL, "bad reference arguments");
    uint8_t *g = nullptr;
    int64_t len = 0;
    int rc = clean_fasta_device(data, size, devices[0], &g, &len);
    if (rc != DRM_OK) return rc;
    std::vector<char> host((size_t)std::max<int64_t>(len, 1));
    cudaError_t e = len ? cudaMemcpy(host.data(), g, (size_t)len, cudaMemcpyDeviceToHost) : cudaSuccess;
    cudaFree(g);
    if (e != cudaSuccess) return fail(DRM_ECUDA, std::string("drm_ref_open_fasta: ") + cudaGetErrorString(e));
    if (genome_len) *genome_len = len;
    return open_ref(host.data(), len, len, 0, ref_len, devices, n_dev, out);
}

int drm_reads_open_fastq(const char *data, int64_t size, const int *devices, int n_dev, drm_reads **out)
{
    if (!out) return fail(DRM_EINVAL, "null output handle");
    *out = nullptr;
    if (!devices || n_dev < 1) return fail(DRM_EINVAL, "bad arguments to drm_reads_open_fastq");
    drm_reads *r = new drm_reads();
    for (int i = 0; i < n_dev; ++i) {
        ReadsRep rep;
        rep.device = devices[i];
        int64_t n = 0, bytes = 0;
        int max_len = 0;
        int rc = parse_fastq_device(data, size, devices[i], &rep.flat, &rep.off, &n, &bytes, &max_len);
        r->rep.push_back(rep);
        if (rc != DRM_OK) { drm_reads_close(r); return rc; }
        r->n = n; r->bytes = bytes; r->max_len = max_len;
    }
    *out = r;
    return DRM_OK;
}

void drm_reads_close(drm_reads *r)
{
    if (!r) return;
    for (auto &rep : r->rep) {
        if (rep.device >= 0) cudaSetDevice(rep.device);
        cudaFree(rep.flat); cudaFree(rep.off);
    }
    delete r;
}

int drm_reads_info(const drm_reads *r, int64_t *info)
{
    if (!r || !info) return fail(DRM_EINVAL, "null argument");
    info[0] = r->n; info[1] = r->bytes; info[2] = r->max_len; info[3] = (int64_t)r->rep.size();
    return DRM_OK;
}

int drm_reads_copy_out(const drm_reads *r, char *flat, int64_t *off)
{
    if (!r || r->rep.empty()) return fail(DRM_EINVAL, "null reads handle");
    DRM_CUDA_TRY(cudaSetDevice(r->rep[0].device));
    if (flat && r->bytes) DRM_CUDA_TRY(cudaMemcpy(flat, r->rep[0].flat, (size_t)r->bytes, cudaMemcpyDeviceToHost));
    if (off) DRM_CUDA_TRY(cudaMemcpy(off, r->rep[0].off, (size_t)(r->n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost));
    return DRM_OK;
}

int drm_reads_device(const drm_reads *r, int slot, const char **flat_dev, const int64_t **off_dev)
{
    if (!r || slot < 0 || slot >= (int)r->rep.size()) return fail(DRM_EINVAL, "bad reads handle or replica slot");
    if (flat_dev) *flat_dev = r->rep[(size_t)slot].flat;
    if (off_dev) *off_dev = r->rep[(size_t)slot].off;
    return DRM_OK;
}

int drm_merge_topk_device(int device, const float *D_all, const int64_t *I_all, int n_shards, int64_t n, int k,
                          float *D_out, int64_t *I_out, void *stream)
{
    DRM_CUDA_TRY(cudaSetDevice(device));
    return launch_merge_topk(D_all, I_all, n_shards, n, k, D_out, I_out, static_cast<cudaStream_t>(stream));
}

/* ------------------------------------------------------------------ embedder */
namespace {

int open_embedder(const EmbedWeightsHost &w, int max_len, int device, drm_embedder **out)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count < 1) return fail(DRM_ECUDA, "no CUDA device: the embedder has no CPU fallback");
    if (device < 0 || device >= count) return fail(DRM_EINVAL, "bad device for the embedder");
    drm_embedder *e = new drm_embedder();
    const int rc = embedder_upload(w, device, max_len, e->dev);
    if (rc != DRM_OK) { embedder_free(e->dev); delete e; return rc; }
    *out = e;
    return DRM_OK;
}

const DeviceRef *ref_on_device(const drm_ref *r, int device)
{
    if (!r) return nullptr;
    for (const DeviceRef &d : r->rep)
        if (d.device == device) return &d;
    return nullptr;
}

// runs a source through the embedder and copies the n x 128 result to the host
int embed_to_host(drm_embedder *e, const EmbedSource &src, int64_t first, int64_t n, float *out)
{
    DevBuf d_out;
    int rc;
    if ((rc = d_out.alloc((size_t)n * 128 * sizeof(float)))) return rc;
    if ((rc = launch_embed(e->dev, src, first, n, d_out.as<float>(), 0)) != DRM_OK) return rc;
    DRM_CUDA_TRY(cudaMemcpy(out, d_out.p, (size_t)n * 128 * sizeof(float), cudaMemcpyDeviceToHost));
    return DRM_OK;
}

int window_source(const drm_embedder *e, const drm_ref *ref, uint64_t first_id, uint64_t id_step, EmbedSource &src)
{
    const DeviceRef *dr = ref_on_device(ref, e->dev.device);
    if (!dr) return fail(DRM_EINVAL, "the reference has no replica on the embedder's device");
    src.mode = 1; src.cls = dr->data; src.form = dr->form; src.ref_n = dr->len; src.ref_len = dr->ref_len;
    src.first_id = first_id; src.id_step = id_step;
    return DRM_OK;
}

}  // namespace

int drm_embedder_open_ir(const char *xml_path, const char *bin_path, int device, drm_embedder **out)
{
    if (!out) return fail(DRM_EINVAL, "null output handle");
    *out = nullptr;
    if (!xml_path || !bin_path) return fail(DRM_EINVAL, "null model path");
    EmbedWeightsHost w;
    const int rc = read_ir_model(xml_path, bin_path, w);
    if (rc != DRM_OK) return rc;
    return open_embedder(w, 123, device, out);
}

int drm_embedder_open_raw(const uint16_t *emb, int vocab, const uint16_t *W1, const uint16_t *R1, const uint16_t *B1,
                          const uint16_t *W2, const uint16_t *R2, const uint16_t *B2, int max_len, int device, drm_embedder **out)
{
    if (!out) return fail(DRM_EINVAL, "null output handle");
    *out = nullptr;
    if (!emb || vocab < 1 || !W1 || !R1 || !B1 || !W2 || !R2 || !B2) return fail(DRM_EINVAL, "bad embedder weights");
    EmbedWeightsHost w;
    w.vocab = vocab;
    auto conv = [](const uint16_t *src, size_t n, std::vector<float> &dst) {
        dst.resize(n);
        for (size_t i = 0; i < n; ++i) dst[i] = half_bits_to_float(src[i]);
    };
    conv(emb, (size_t)vocab * 64, w.emb);
    conv(W1, 2u * 192 * 64, w.W1); conv(R1, 2u * 192 * 64, w.R1); conv(B1, 2u * 256, w.B1);
    conv(W2, 2u * 192 * 128, w.W2); conv(R2, 2u * 192 * 64, w.R2); conv(B2, 2u * 256, w.B2);
    return open_embedder(w, max_len, device, out);
}

void drm_embedder_close(drm_embedder *e)
{
    if (!e) return;
    if (e->wcache.emb) {
        if (e->dev.device >= 0) cudaSetDevice(e->dev.device);
        cudaFree(e->wcache.emb);
    }
    embedder_free(e->dev);
    delete e;
}

int drm_embedder_ir_info(const char *xml_path, const char *bin_path, int64_t *info, double *sums)
{
    if (!xml_path || !bin_path || !info) return fail(DRM_EINVAL, "null argument");
    EmbedWeightsHost w;
    const int rc = read_ir_model(xml_path, bin_path, w);
    if (rc != DRM_OK) return rc;
    const std::vector<float> *t[7] = {&w.emb, &w.W1, &w.R1, &w.B1, &w.W2, &w.R2, &w.B2};
    info[0] = w.vocab;
    for (int i = 0; i < 7; ++i) {
        info[1 + i] = (int64_t)t[i]->size();
        if (sums) {
            double s = 0.0;
            for (float v : *t[i]) s += v;
            sums[i] = s;
        }
    }
    return DRM_OK;
}

int drm_tokenize(const char *seq_flat, const int64_t *seq_off, int64_t n, int max_len, uint16_t *tokens)
{
    if (n <= 0) return DRM_OK;
    if (!seq_flat || !seq_off || !tokens || max_len < 2 || max_len > 128) return fail(DRM_EINVAL, "bad arguments to drm_tokenize");
    tokenize_host(seq_flat, seq_off, n, max_len, tokens);
    return DRM_OK;
}

int drm_embed(drm_embedder *e, const char *seq_flat, const int64_t *seq_off, int64_t n, float *out)
{
    if (!e) return fail(DRM_EINVAL, "null embedder handle");
    if (n <= 0) return DRM_OK;
    if (!seq_flat || !seq_off || !out) return fail(DRM_EINVAL, "bad arguments to drm_embed");
    DRM_CUDA_TRY(cudaSetDevice(e->dev.device));
    const int64_t bytes = seq_off[n] - seq_off[0];
    std::vector<int64_t> rel((size_t)n + 1);
    for (int64_t i = 0; i <= n; ++i) {
        rel[(size_t)i] = seq_off[i] - seq_off[0];
        if (i && rel[(size_t)i] < rel[(size_t)i - 1]) return fail(DRM_EINVAL, "sequence offsets must not decrease");
    }
    DevBuf d_flat, d_off;
    int rc;
    if ((rc = d_flat.alloc((size_t)bytes)) || (rc = d_off.alloc((size_t)(n + 1) * 8))) return rc;
    if (bytes) DRM_CUDA_TRY(cudaMemcpy(d_flat.p, seq_flat + seq_off[0], (size_t)bytes, cudaMemcpyHostToDevice));
    DRM_CUDA_TRY(cudaMemcpy(d_off.p, rel.data(), (size_t)(n + 1) * 8, cudaMemcpyHostToDevice));
    EmbedSource src;
    src.mode = 0; src.flat = d_flat.as<char>(); src.off = d_off.as<int64_t>();
    return embed_to_host(e, src, 0, n, out);
}

int drm_embed_reads(drm_embedder *e, const drm_reads *reads, int64_t first, int64_t n, float *out)
{
    if (!e || !reads) return fail(DRM_EINVAL, "null handle");
    if (n <= 0) return DRM_OK;
    if (!out || first < 0 || first + n > reads->n) return fail(DRM_EINVAL, "bad read range for drm_embed_reads");
    const ReadsRep *rep = nullptr;
    for (const ReadsRep &r : reads->rep)
        if (r.device == e->dev.device) rep = &r;
    if (!rep) return fail(DRM_EINVAL, "the reads have no replica on the embedder's device");
    DRM_CUDA_TRY(cudaSetDevice(e->dev.device));
    EmbedSource src;
    src.mode = 0; src.flat = rep->flat; src.off = rep->off;
    return embed_to_host(e, src, first, n, out);
}

int drm_embed_windows(drm_embedder *e, const drm_ref *ref, uint64_t first_id, uint64_t id_step, int64_t n, float *out)
{
    if (!e || !ref) return fail(DRM_EINVAL, "null handle");
    if (n <= 0) return DRM_OK;
    if (!out) return fail(DRM_EINVAL, "null output");
    EmbedSource src;
    int rc = window_source(e, ref, first_id, id_step, src);
    if (rc != DRM_OK) return rc;
    DRM_CUDA_TRY(cudaSetDevice(e->dev.device));
    return embed_to_host(e, src, 0, n, out);
}

int drm_embed_device(drm_embedder *e, const char *seq_flat_dev, const int64_t *seq_off_dev, int64_t first, int64_t n,
                     float *out_dev, void *stream)
{
    if (!e) return fail(DRM_EINVAL, "null embedder handle");
    if (n <= 0) return DRM_OK;
    if (!seq_flat_dev || !seq_off_dev || !out_dev || first < 0) return fail(DRM_EINVAL, "bad arguments to drm_embed_device");
    DRM_CUDA_TRY(cudaSetDevice(e->dev.device));
    EmbedSource src;
    src.mode = 0; src.flat = seq_flat_dev; src.off = seq_off_dev;
    return launch_embed(e->dev, src, first, n, out_dev, static_cast<cudaStream_t>(stream));
}

int drm_embed_windows_device(drm_embedder *e, const drm_ref *ref, uint64_t first_id, uint64_t id_step, int64_t n,
                             float *out_dev, void *stream)
{
    if (!e || !ref) return fail(DRM_EINVAL, "null handle");
    if (n <= 0) return DRM_OK;
    if (!out_dev) return fail(DRM_EINVAL, "null output");
    EmbedSource src;
    int rc = window_source(e, ref, first_id, id_step, src);
    if (rc != DRM_OK) return rc;
    DRM_CUDA_TRY(cudaSetDevice(e->dev.device));
    return launch_embed(e->dev, src, 0, n, out_dev, static_cast<cudaStream_t>(stream));
}

/* ------------------------------------------------------------------ L2 rerank stage */
int drm_l2_plan_host(int mode, const int64_t *neighbors, int64_t n, int64_t k_in, int64_t stride, int64_t k, int64_t k_clusters,
                     uint64_t bound, int64_t *info, uint64_t *unique, uint32_t *map, int64_t *starts)
{
    if (!neighbors || !info) return fail(DRM_EINVAL, "null argument");
    L2Plan plan;
    // a negative mode asks for the plan drm_l2_rerank makes when every window's embedding is at hand: -1 - mode, window ids in `map`
    const bool direct = mode < 0;
    const int rc = l2_plan(direct ? -1 - mode : mode, neighbors, n, k_in, stride, k, k_clusters, bound, plan, direct);
    if (rc != DRM_OK) return rc;
    info[0] = (int64_t)plan.unique.size(); info[1] = (int64_t)plan.map.size(); info[2] = plan.k_out;
    info[3] = plan.passthrough ? 1 : 0; info[4] = plan.take;
    if (unique) std::copy(plan.unique.begin(), plan.unique.end(), unique);
    if (map) std::copy(plan.map.begin(), plan.map.end(), map);
    if (starts) std::copy(plan.starts.begin(), plan.starts.end(), starts);
    return DRM_OK;
}

int drm_l2_select_host(const float *dist, int64_t n, int64_t k, int64_t *idx)
{
    if (!dist || !idx || n < 0 || k < 0) return fail(DRM_EINVAL, "bad arguments to drm_l2_select_host");
    l2_select_host(dist, n, k, idx);
    return DRM_OK;
}

static thread_local double g_l2_ms[5] = {0, 0, 0, 0, 0};  // plan (host), embed, distance + selection, copies + bookkeeping, batches

int drm_l2_last_timing(double *ms5)
{
    if (!ms5) return fail(DRM_EINVAL, "null output");
    for (int i = 0; i < 5; ++i) ms5[i] = g_l2_ms[i];
    return DRM_OK;
}

int drm_l2_rerank(drm_embedder *e, drm_ref *ref, int mode, const int64_t *neighbors, const float *distances, int64_t n, int64_t k_in,
                  const float *q_emb, int64_t stride, int64_t k, int64_t k_clusters, float *out_scores, uint64_t *out_ids,
                  int64_t *out_count)
{
    if (!e || !ref) return fail(DRM_EINVAL, "null handle");
    if (!out_count) return fail(DRM_EINVAL, "null count output");
    *out_count = 0;
    if (n <= 0) return DRM_OK;
    if (!neighbors || !distances || !q_emb || !out_scores || !out_ids || (mode < 0 || mode > 2)) return fail(DRM_EINVAL, "bad arguments to drm_l2_rerank");
    const DeviceRef *dr = ref_on_device(ref, e->dev.device);
    if (!dr) return fail(DRM_EINVAL, "the reference has no replica on the embedder's device");
    if (mode == 1 && dr->form != 0) return fail(DRM_EINVAL, "dynamic mode needs a genome reference (drm_ref_open_genome)");
    // bound: dynamic = genome length in bases; static = number of ref_seqs (rows of the window table, or the dense windows of the genome)
    uint64_t bound;
    if (mode == 1 || dr->form == 1) bound = (uint64_t)dr->len;
    else bound = dr->len >= dr->ref_len ? 2ull * (uint64_t)(dr->len - dr->ref_len + 1) : 0;
    L2Plan plan;
    for (double &v : g_l2_ms) v = 0;
    // Window cache: a window's embedding depends on its id alone, and every candidate id lies below `bound`.  Once a call's flat
    // candidate list is as long as the reference has windows (one E. coli batch of 20 000 reads lists twice that), embedding ALL
    // `bound` windows once - 0.5 s and 2.4 GB for E. coli - is cheaper than embedding each batch's share, and later calls on the
    // same reference embed nothing at all: the kernels read candidate rows straight from the cache by window id, and the host
    // plan skips its de-duplication pass.  Same values as the per-batch path (the embedder treats sequences independently).
    // DRM_L2_CACHE_GB = budget (default 16, 0 = off); DRM_L2_CACHE_SHARE = flat candidates per window that trigger the build
    // (default 1; tests: 0 = always).
    double cache_gb = 16.0, share = 1.0;
    if (const char *ev = getenv("DRM_L2_CACHE_GB")) cache_gb = atof(ev);
    if (const char *ev = getenv("DRM_L2_CACHE_SHARE")) share = atof(ev);
    const int decorate = mode == 0 ? 1 : 0;  // mode 0: ref_seqs hold '<' window '>' strings; find_sequence (1) and lookup-mode tables (2) give bare windows
    drm_embedder::WindowCache &wc = e->wcache;
    const bool cache_fits = bound > 0 && bound < 0xffffffffull && (double)bound * 512.0 <= cache_gb * 1073741824.0;
    const bool cache_hit = cache_fits && wc.emb && wc.ref_uid == ref->uid && wc.decorate == decorate && wc.n == bound;
    const double flat_est = (double)n * (double)std::min(mode == 1 ? k_clusters : k_in, k_in) * (double)(mode == 1 ? 2 * stride - 1 : stride);
    bool use_cache = cache_hit || (cache_fits && flat_est >= share * (double)bound && !(mode == 1 && stride == 1));
    if (use_cache && !cache_hit) {  // room for the table first: without it the call takes the per-batch path, as before
        DRM_CUDA_TRY(cudaSetDevice(e->dev.device));
        if (wc.emb) { cudaFree(wc.emb); wc = drm_embedder::WindowCache(); }
        float *fresh = nullptr;
        if (cudaMalloc(reinterpret_cast<void **>(&fresh), (size_t)bound * 128 * sizeof(float)) != cudaSuccess) {
            cudaGetLastError();
            use_cache = false;
        } else wc.emb = fresh;  // owned by the handle from here on; its key is set once it is filled
    }
    const auto t_plan = std::chrono::steady_clock::now();
    int rc = l2_plan(mode == 2 ? 0 : mode, neighbors, n, k_in, stride, k, k_clusters, bound, plan, use_cache);  // mode 2 = static bookkeeping, bare windows
    if (rc != DRM_OK) return rc;
    g_l2_ms[0] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_plan).count();
    const auto t_all = std::chrono::steady_clock::now();
    cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_c = nullptr;
    if (plan.passthrough) {
        for (int64_t q = 0; q < n; ++q)
            for (int64_t j = 0; j < plan.take; ++j) {
                out_scores[*out_count] = distances[q * k_in + j];
                out_ids[*out_count] = (uint64_t)neighbors[q * k_in + j];
                ++*out_count;
            }
        return DRM_OK;
    }
    const int64_t U = (int64_t)plan.unique.size(), ko = plan.k_out;
    if (ko < 1) return fail(DRM_EINVAL, "k must be >= 1");
    DRM_CUDA_TRY(cudaSetDevice(e->dev.device));
    DRM_CUDA_TRY(cudaEventCreate(&ev_a)); DRM_CUDA_TRY(cudaEventCreate(&ev_b)); DRM_CUDA_TRY(cudaEventCreate(&ev_c));
    struct EvGuard { cudaEvent_t &a, &b, &c; ~EvGuard() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); if (c) cudaEventDestroy(c); } } ev_guard{ev_a, ev_b, ev_c};
    // The plan (the reference's one flat candidate list for ALL queries, de-duplicated windows) stays on the host; the device
    // works through it in QUERY BATCHES bounded by a memory budget, like post_process_l2_dynamic_streaming batches its queries:
    // a batch embeds only the windows its own slices reference (a window's embedding does not depend on who else asks for it,
    // so the result is the same) and reuses the buffers of the previous batch.  Budget: DRM_L2_BATCH_CANDS flat candidates per
    // batch (default 8 M: <= 4 GB of embeddings + 0.1 GB of scratch).
    int64_t max_flat = 8ll << 20;
    if (const char *ev = getenv("DRM_L2_BATCH_CANDS")) max_flat = std::max<int64_t>(1, atoll(ev));
    if (use_cache && !cache_hit) {
        EmbedSource all;
        all.mode = 1; all.cls = dr->data; all.form = dr->form; all.ref_n = dr->len; all.ref_len = dr->ref_len;
        all.first_id = 0; all.id_step = 1; all.ids = nullptr; all.decorate = decorate;
        DRM_CUDA_TRY(cudaEventRecord(ev_a, 0));
        if ((rc = launch_embed(e->dev, all, 0, (int64_t)bound, wc.emb, 0)) != DRM_OK) { cudaFree(wc.emb); wc = drm_embedder::WindowCache(); return rc; }
        DRM_CUDA_TRY(cudaEventRecord(ev_b, 0));
        DRM_CUDA_TRY(cudaEventSynchronize(ev_b));
        float mb = 0;
        cudaEventElapsedTime(&mb, ev_a, ev_b);
        g_l2_ms[1] += mb;
        wc.ref_uid = ref->uid; wc.decorate = decorate; wc.n = bound;
    }
    std::vector<int32_t> local_of((size_t)U, -1);   // global unique row -> row of this batch's embedding table
    std::vector<uint64_t> b_ids;
    std::vector<uint32_t> b_map, b_rows;
    std::vector<int64_t> b_starts;
    DevBuf d_ids, d_map, d_starts, d_q, d_emb, d_dist, d_idx, d_od, d_or, d_cnt;
    size_t cap_u = 0, cap_f = 0, cap_n = 0;
    std::vector<float> od;
    std::vector<uint32_t> orow;
    std::vector<int64_t> cnt;
    for (int64_t q0 = 0; q0 < n;) {
        int64_t q1 = q0 + 1;
        while (q1 < n && plan.starts[(size_t)q1 + 1] - plan.starts[(size_t)q0] <= max_flat) ++q1;
        const int64_t nb = q1 - q0, f0 = plan.starts[(size_t)q0], fb = plan.starts[(size_t)q1] - f0;
        b_ids.clear(); b_rows.clear();
        const bool map_in_place = use_cache && plan.direct_ids;  // the plan's own flat list is what the device needs
        if (!map_in_place) b_map.resize((size_t)fb);
        b_starts.resize((size_t)nb + 1);
        for (int64_t i = 0; i <= nb; ++i) b_starts[(size_t)i] = plan.starts[(size_t)(q0 + i)] - f0;
        if (use_cache) {  // candidate rows = window ids into the cache: no per-batch table, nothing to embed
            if (!plan.direct_ids)
                for (int64_t j = 0; j < fb; ++j) b_map[(size_t)j] = (uint32_t)plan.unique[plan.map[(size_t)(f0 + j)]];
        } else {
            for (int64_t j = 0; j < fb; ++j) {
                const uint32_t g = plan.map[(size_t)(f0 + j)];
                int32_t l = local_of[g];
                if (l < 0) { l = (int32_t)b_ids.size(); local_of[g] = l; b_ids.push_back(plan.unique[g]); b_rows.push_back(g); }
                b_map[(size_t)j] = (uint32_t)l;
            }
            for (uint32_t g : b_rows) local_of[g] = -1;
        }
        const int64_t ub = (int64_t)b_ids.size();
        if ((size_t)ub > cap_u || (size_t)fb > cap_f || (size_t)nb > cap_n) {  // grow (first batch sizes the rest: batches are cut to one budget)
            cap_u = std::max(cap_u, (size_t)ub + (size_t)ub / 8 + 16); cap_f = std::max(cap_f, (size_t)fb + 16); cap_n = std::max(cap_n, (size_t)nb + 16);
            DRM_CUDA_TRY(cudaDeviceSynchronize());
            d_ids.release(); d_map.release(); d_starts.release(); d_q.release(); d_emb.release(); d_dist.release(); d_idx.release(); d_od.release();
            d_or.release(); d_cnt.release();
            if ((rc = d_ids.alloc(cap_u * 8)) || (rc = d_map.alloc(cap_f * 4)) || (rc = d_starts.alloc((cap_n + 1) * 8)) || (rc = d_q.alloc(cap_n * 128 * 4)) ||
                (rc = d_emb.alloc(cap_u * 128 * 4)) || (rc = d_dist.alloc(cap_f * 4)) || (rc = d_idx.alloc(cap_f * 4)) || (rc = d_od.alloc(cap_n * ko * 4)) ||
                (rc = d_or.alloc(cap_n * ko * 4)) || (rc = d_cnt.alloc(cap_n * 8))) return rc;
        }
        if (ub) DRM_CUDA_TRY(cudaMemcpy(d_ids.p, b_ids.data(), (size_t)ub * 8, cudaMemcpyHostToDevice));
        if (fb) DRM_CUDA_TRY(cudaMemcpy(d_map.p, map_in_place ? plan.map.data() + f0 : b_map.data(), (size_t)fb * 4, cudaMemcpyHostToDevice));
        DRM_CUDA_TRY(cudaMemcpy(d_starts.p, b_starts.data(), (size_t)(nb + 1) * 8, cudaMemcpyHostToDevice));
        DRM_CUDA_TRY(cudaMemcpy(d_q.p, q_emb + (size_t)q0 * 128, (size_t)nb * 128 * 4, cudaMemcpyHostToDevice));
        EmbedSource src;
        src.mode = 1; src.cls = dr->data; src.form = dr->form; src.ref_n = dr->len; src.ref_len = dr->ref_len;
        src.ids = d_ids.as<unsigned long long>();
        src.decorate = decorate;
        DRM_CUDA_TRY(cudaEventRecord(ev_a, 0));
        if (ub && (rc = launch_embed(e->dev, src, 0, ub, d_emb.as<float>(), 0)) != DRM_OK) return rc;
        DRM_CUDA_TRY(cudaEventRecord(ev_b, 0));
        if ((rc = launch_l2_rerank(use_cache ? wc.emb : d_emb.as<float>(), d_map.as<uint32_t>(), d_starts.as<int64_t>(), d_q.as<float>(), nb, (int)ko, d_dist.as<float>(),
                                   d_idx.as<int64_t>(), d_od.as<float>(), d_or.as<uint32_t>(), d_cnt.as<int64_t>(), 0)) != DRM_OK) return rc;
        DRM_CUDA_TRY(cudaEventRecord(ev_c, 0));
        od.resize((size_t)(nb * ko)); orow.resize((size_t)(nb * ko)); cnt.resize((size_t)nb);
        DRM_CUDA_TRY(cudaMemcpy(od.data(), d_od.p, od.size() * 4, cudaMemcpyDeviceToHost));
        DRM_CUDA_TRY(cudaMemcpy(orow.data(), d_or.p, orow.size() * 4, cudaMemcpyDeviceToHost));
        DRM_CUDA_TRY(cudaMemcpy(cnt.data(), d_cnt.p, cnt.size() * 8, cudaMemcpyDeviceToHost));
        for (int64_t q = 0; q < nb; ++q) {
            if (cnt[(size_t)q] == 0) continue;  // batch_reranker returns empty vectors for a query without candidates
            if (cnt[(size_t)q] < 0)
                return fail(DRM_ENOTENOUGH, "Not enough candidates (" + std::to_string(-cnt[(size_t)q] - 1) + " < " + std::to_string(ko) + ") for query " + std::to_string(q0 + q));
            for (int64_t j = 0; j < ko; ++j) {
                out_scores[*out_count] = od[(size_t)(q * ko + j)];
                out_ids[*out_count] = use_cache ? (uint64_t)orow[(size_t)(q * ko + j)] : b_ids[orow[(size_t)(q * ko + j)]];
                ++*out_count;
            }
        }
        {
            float m1 = 0, m2 = 0;
            cudaEventElapsedTime(&m1, ev_a, ev_b);
            cudaEventElapsedTime(&m2, ev_b, ev_c);
            g_l2_ms[1] += m1; g_l2_ms[2] += m2; g_l2_ms[4] += 1;
        }
        q0 = q1;
    }
    g_l2_ms[3] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_all).count() - g_l2_ms[1] - g_l2_ms[2];
    return DRM_OK;
}

}  // extern "C"
