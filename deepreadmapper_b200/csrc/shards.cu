// Position-sharded index behind the C ABI (BASELINE.json configs[4], SURVEY.md §8b `shard_mode = 1`, §8e): the reference
// windows are cut by position into W shards, one HNSW-PQ graph per GPU; every read is searched on every shard and the
// per-shard top-k lists meet on the GPU that owns the read (reads are dealt out in W contiguous slices per chunk), which
// merges them (merge_kernel.cu) and - in the *_rerank entry point - SW-reranks its slice from the replicated genome.
// Replaces, for an index too large for one GPU, what the reference does with one faiss_search call on one index
// (src/main.cpp:236-237,270-296).
//
// Three exchange modes, same results bit for bit (tests/test_gpu_shards.py):
//   DRM_XCHG_ALLGATHER (0)  ncclAllGather of every rank's (D, I)[chunk][k], each rank then merges its slice.  The
//                           north_star's baseline: W x chunk x k x 12 bytes land on EVERY GPU.
//   DRM_XCHG_ALLTOALL  (1)  grouped ncclSend / ncclRecv: a rank receives only the lists of the reads it merges -
//                           chunk x k x 12 bytes per GPU, 1/W of the all-gather.
//   DRM_XCHG_FUSED     (2)  no exchange step at all: the search kernel's epilogue (search_exact.cu, SearchDispatch) writes
//                           each read's list - labels already global - straight into the owner's merge buffer, local
//                           memory or a peer GPU's over NVLink (cudaDeviceEnablePeerAccess in one process, cudaIpc
//                           handles between processes), list by list while the rest of the chunk is still being searched.
//                           What remains of the "collective" is one 4-byte ncclAllReduce per chunk as the barrier that
//                           orders the peers' writes before the merge.
// Two process models: one process driving all W GPUs (drm_shards_open: one host thread per GPU; what `pipeline` uses under
// DRM_SHARD=position) and one process per GPU (drm_shards_open_rank: torchrun / mpirun; the ncclUniqueId travels through
// whatever the launcher offers).  NCCL is loaded with dlopen at first use, so libdrm_cuda.so has no link-time dependency
// on it and every non-sharded entry point works on a box without NCCL.
#include "common.h"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>
#include <thread>

namespace drm {

namespace {

struct NcclApi {
    void *h = nullptr;
    bool tried = false;
    std::string err;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
};

NcclApi &nccl()
{
    static NcclApi a;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    if (a.tried) return a;
    a.tried = true;
    // An NCCL that the process has loaded already (a PyTorch process carries its own, newer one) wins: two different
    // libnccl.so.2 in one process do not mix.  Otherwise DRM_NCCL_LIB, then the system library.
    a.h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL | RTLD_NOLOAD);
    const char *names[] = {getenv("DRM_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
        if (a.h) break;
        if (!nm || !*nm) continue;
        a.h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (!a.h) a.err = dlerror();
    }
    if (!a.h) { a.err = "NCCL not found (" + a.err + "); set DRM_NCCL_LIB"; return a; }
#define LOAD(field, sym)                                                             \
    *reinterpret_cast<void **>(&a.field) = dlsym(a.h, sym);                          \
    if (!a.field) { a.err = std::string("NCCL symbol missing: ") + sym; a.h = nullptr; return a; }
    LOAD(GetUniqueId, "ncclGetUniqueId") LOAD(CommInitRank, "ncclCommInitRank") LOAD(CommInitAll, "ncclCommInitAll")
    LOAD(CommDestroy, "ncclCommDestroy") LOAD(AllGather, "ncclAllGather") LOAD(AllReduce, "ncclAllReduce") LOAD(Send, "ncclSend")
    LOAD(Recv, "ncclRecv") LOAD(GroupStart, "ncclGroupStart") LOAD(GroupEnd, "ncclGroupEnd") LOAD(GetErrorString, "ncclGetErrorString")
    LOAD(GetVersion, "ncclGetVersion")
#undef LOAD
    return a;
}

#define DRM_NCCL_TRY(expr)                                                                                       \
    do {                                                                                                         \
        ncclResult_t r__ = (expr);                                                                               \
        if (r__ != ncclSuccess) return ::drm::fail(DRM_ECUDA, std::string(#expr) + ": " + nccl().GetErrorString(r__)); \
    } while (0)

struct ShardRank {
    int device = -1, rank = 0, world = 1;
    int64_t id_lo = 0;
    DeviceIndex dx;
    ncclComm_t comm = nullptr;
    cudaStream_t s = nullptr, s_copy = nullptr;
    // capacities the buffers below were sized for
    int64_t cap_chunk = 0, slice_cap = 0;
    int cap_k = 0, d = 0, cap_mode = -1;
    float *q[2] = {nullptr, nullptr};
    cudaEvent_t q_ready[2] = {nullptr, nullptr}, q_free[2] = {nullptr, nullptr};
    float *D_loc = nullptr; int64_t *I_loc = nullptr;                            // [chunk][k]        modes 0, 1
    float *D_ag = nullptr; int64_t *I_ag = nullptr;                              // [W][chunk][k]     mode 0
    float *D_rx[2] = {nullptr, nullptr}; int64_t *I_rx[2] = {nullptr, nullptr};  // [W][slice_cap][k] modes 1, 2 (2: P2P targets)
    float *D_m[2] = {nullptr, nullptr}; int64_t *I_m[2] = {nullptr, nullptr};    // [slice_cap][k]    merged lists of my slice
    int *flag = nullptr;
    float *D_peer[2][MAX_SHARDS]; int64_t *I_peer[2][MAX_SHARDS];                // every rank's rx buffers seen from here
    std::vector<void *> ipc_open;                                                // cudaIpcOpenMemHandle results to close
    bool p2p_ready = false;
    int64_t *counters = nullptr;                                                 // nhops0, nhops_up, ndis (device)
    // rerank staging of the host-buffer call: two sets, so that chunk c + 1 is computed while chunk c's results go home
    char *rq[2] = {nullptr, nullptr}; int64_t *rqo[2] = {nullptr, nullptr}; int32_t *rsc[2] = {nullptr, nullptr}; uint64_t *rid[2] = {nullptr, nullptr};
    int64_t *rel_h[2] = {nullptr, nullptr};                                      // pinned host: chunk-relative read offsets
    int32_t *rmin = nullptr;
    size_t cap_rq = 0, cap_rqo = 0, cap_rout = 0;
    cudaStream_t s_out = nullptr;                                                // device -> host copies of finished slices
    cudaEvent_t out_done[2] = {nullptr, nullptr}, rel_copied[2] = {nullptr, nullptr}, res_ready[2] = {nullptr, nullptr};
    int64_t step = 0;
    // stage boundaries of the last EV_SETS steps (search | exchange | merge | rerank); harvested lazily into the sums
    static constexpr int EV_SETS = 8;
    cudaEvent_t evr[EV_SETS][5] = {};
    cudaEvent_t *ev = evr[0];
    bool ev_pending[EV_SETS] = {};
    int ev_last[EV_SETS] = {};
    int ev_cur = 0;
    double ms_search = 0, ms_xchg = 0, ms_merge = 0, ms_rerank = 0;
    int64_t bytes_rx = 0;
};

void free_rank_buffers(ShardRank &r)
{
    if (r.device >= 0) cudaSetDevice(r.device);
    for (void *p : r.ipc_open) cudaIpcCloseMemHandle(p);
    r.ipc_open.clear();
    for (int b = 0; b < 2; ++b) {
        cudaFree(r.q[b]); cudaFree(r.D_rx[b]); cudaFree(r.I_rx[b]); cudaFree(r.D_m[b]); cudaFree(r.I_m[b]);
        r.q[b] = nullptr; r.D_rx[b] = nullptr; r.I_rx[b] = nullptr; r.D_m[b] = nullptr; r.I_m[b] = nullptr;
    }
    cudaFree(r.D_loc); cudaFree(r.I_loc); cudaFree(r.D_ag); cudaFree(r.I_ag);
    r.D_loc = nullptr; r.I_loc = nullptr; r.D_ag = nullptr; r.I_ag = nullptr;
    r.cap_chunk = 0; r.cap_k = 0; r.cap_mode = -1; r.p2p_ready = false;
}

void harvest_stage_times(ShardRank &r, int set)
{
    if (!r.ev_pending[set]) return;
    r.ev_pending[set] = false;
    cudaEvent_t *e = r.evr[set];
    if (cudaEventSynchronize(e[r.ev_last[set]]) != cudaSuccess) { cudaGetLastError(); return; }
    float a = 0, b = 0, c = 0, d = 0;
    cudaEventElapsedTime(&a, e[0], e[1]);
    cudaEventElapsedTime(&b, e[1], e[2]);
    cudaEventElapsedTime(&c, e[2], e[3]);
    if (r.ev_last[set] == 4) cudaEventElapsedTime(&d, e[3], e[4]);
    r.ms_search += a; r.ms_xchg += b; r.ms_merge += c; r.ms_rerank += d;
}

void harvest_all(ShardRank &r)
{
    for (int i = 0; i < ShardRank::EV_SETS; ++i) harvest_stage_times(r, i);
}

}  // namespace
}  // namespace drm

struct drm_shards {
    std::vector<std::unique_ptr<drm::ShardRank>> local;  // W ranks in the one-process model, 1 in the rank-per-process model
    int world = 1;
    bool one_process = true;
    bool peer_ok = false;  // every pair of GPUs can map each other's memory
    std::mutex mu;
};

namespace drm {
namespace {

int ensure_buffers(drm_shards *h, ShardRank &r, int64_t chunk, int k, int d, int mode)
{
    const int W = r.world;
    if (r.cap_chunk >= chunk && r.cap_k == k && r.d == d && r.cap_mode == mode) return DRM_OK;
    // (re)allocation is collective: every rank gets here with the same arguments
    DRM_CUDA_TRY(cudaSetDevice(r.device));
    DRM_CUDA_TRY(cudaStreamSynchronize(r.s));
    free_rank_buffers(r);
    const int64_t slice_cap = (chunk + W - 1) / W + 1;
    const size_t lk = (size_t)chunk * k, sk = (size_t)slice_cap * k;
    for (int b = 0; b < 2; ++b) {
        DRM_CUDA_TRY(cudaMalloc(&r.q[b], (size_t)chunk * d * sizeof(float)));
        DRM_CUDA_TRY(cudaMalloc(&r.D_m[b], sk * sizeof(float)));
        DRM_CUDA_TRY(cudaMalloc(&r.I_m[b], sk * sizeof(int64_t)));
        if (W > 1 && mode != DRM_XCHG_ALLGATHER) {
            DRM_CUDA_TRY(cudaMalloc(&r.D_rx[b], (size_t)W * sk * sizeof(float)));
            DRM_CUDA_TRY(cudaMalloc(&r.I_rx[b], (size_t)W * sk * sizeof(int64_t)));
        }
    }
    if (W == 1 || mode != DRM_XCHG_FUSED) {
        DRM_CUDA_TRY(cudaMalloc(&r.D_loc, lk * sizeof(float)));
        DRM_CUDA_TRY(cudaMalloc(&r.I_loc, lk * sizeof(int64_t)));
    }
    if (W > 1 && mode == DRM_XCHG_ALLGATHER) {
        DRM_CUDA_TRY(cudaMalloc(&r.D_ag, (size_t)W * lk * sizeof(float)));
        DRM_CUDA_TRY(cudaMalloc(&r.I_ag, (size_t)W * lk * sizeof(int64_t)));
    }
    r.cap_chunk = chunk; r.cap_k = k; r.d = d; r.cap_mode = mode; r.slice_cap = slice_cap;
    (void)h;
    return DRM_OK;
}

// Fused mode needs every rank's rx buffers addressable from every device.  One process: the pointers themselves (peer access
// enabled at open).  One process per GPU: cudaIpc handles, all-gathered through NCCL as bytes.
int publish_peers(drm_shards *h, ShardRank &r)
{
    if (r.p2p_ready) return DRM_OK;
    const int W = r.world;
    DRM_CUDA_TRY(cudaSetDevice(r.device));
    if (h->one_process) {
        for (int b = 0; b < 2; ++b)
            for (int o = 0; o < W; ++o) { r.D_peer[b][o] = h->local[(size_t)o]->D_rx[b]; r.I_peer[b][o] = h->local[(size_t)o]->I_rx[b]; }
        r.p2p_ready = true;
        return DRM_OK;
    }
    struct Handles { cudaIpcMemHandle_t m[4]; };
    Handles mine;
    void *ptrs[4] = {r.D_rx[0], r.I_rx[0], r.D_rx[1], r.I_rx[1]};
    for (int i = 0; i < 4; ++i) DRM_CUDA_TRY(cudaIpcGetMemHandle(&mine.m[i], ptrs[i]));
    Handles *d_all = nullptr, *d_mine = nullptr;
    DRM_CUDA_TRY(cudaMalloc(&d_all, sizeof(Handles) * W));
    DRM_CUDA_TRY(cudaMalloc(&d_mine, sizeof(Handles)));
    DRM_CUDA_TRY(cudaMemcpyAsync(d_mine, &mine, sizeof(Handles), cudaMemcpyHostToDevice, r.s));
    DRM_NCCL_TRY(nccl().AllGather(d_mine, d_all, sizeof(Handles), ncclChar, r.comm, r.s));
    std::vector<Handles> all((size_t)W);
    DRM_CUDA_TRY(cudaMemcpyAsync(all.data(), d_all, sizeof(Handles) * W, cudaMemcpyDeviceToHost, r.s));
    DRM_CUDA_TRY(cudaStreamSynchronize(r.s));
    cudaFree(d_all); cudaFree(d_mine);
    for (int o = 0; o < W; ++o) {
        void *p[4];
        if (o == r.rank) { for (int i = 0; i < 4; ++i) p[i] = ptrs[i]; }
        else {
            for (int i = 0; i < 4; ++i) {
                DRM_CUDA_TRY(cudaIpcOpenMemHandle(&p[i], all[(size_t)o].m[i], cudaIpcMemLazyEnablePeerAccess));
                r.ipc_open.push_back(p[i]);
            }
        }
        r.D_peer[0][o] = (float *)p[0]; r.I_peer[0][o] = (int64_t *)p[1]; r.D_peer[1][o] = (float *)p[2]; r.I_peer[1][o] = (int64_t *)p[3];
    }
    r.p2p_ready = true;
    return DRM_OK;
}

inline int64_t slice_lo(int64_t n, int o, int W) { return n * o / W; }

// One chunk on one rank: search every read of the chunk on this shard, exchange, merge this rank's slice.  Everything is
// enqueued on r.s; on return *D_out / *I_out are the merged lists of reads [lo, lo + cnt) of the chunk (device memory of
// the handle, valid until the same slot is used again two steps later).
int rank_step(drm_shards *h, ShardRank &r, const float *q_dev, int64_t n, int k, int ef, int mode, const float **D_out, const int64_t **I_out,
              int64_t *lo_out, int64_t *cnt_out)
{
    const int W = r.world;
    if (W > 1 && !r.comm) return fail(DRM_EINVAL, "sharded handle has no communicator");
    if (!search_exact_supports(r.dx, k, ef)) return fail(DRM_EUNSUPPORTED, "sharded search needs a shape the exact kernel supports (k <= 256)");
    if (mode == DRM_XCHG_FUSED && W > 1 && !h->peer_ok) mode = DRM_XCHG_ALLTOALL;  // no peer mapping between some pair of GPUs
    DRM_CUDA_TRY(cudaSetDevice(r.device));
    int rc = ensure_buffers(h, r, std::max<int64_t>(n, r.cap_chunk), k, r.dx.d, mode);
    if (rc != DRM_OK) return rc;
    const int slot = (int)(r.step & 1);
    r.ev_cur = (int)(r.step % ShardRank::EV_SETS);
    harvest_stage_times(r, r.ev_cur);  // the set used EV_SETS steps ago: long finished
    r.ev = r.evr[r.ev_cur];
    r.ev_pending[r.ev_cur] = true;
    r.ev_last[r.ev_cur] = 3;
    ++r.step;
    const int64_t lo = slice_lo(n, r.rank, W), cnt = slice_lo(n, r.rank + 1, W) - lo;
    *lo_out = lo; *cnt_out = cnt;
    DRM_CUDA_TRY(cudaEventRecord(r.ev[0], r.s));
    if (W == 1) {
        rc = launch_search_exact_ex(r.dx, q_dev, n, k, ef, r.D_m[slot], r.I_m[slot], r.counters, r.s, r.id_lo, nullptr);
        if (rc != DRM_OK) return rc;
        for (int e = 1; e <= 3; ++e) DRM_CUDA_TRY(cudaEventRecord(r.ev[e], r.s));
        *D_out = r.D_m[slot]; *I_out = r.I_m[slot];
        return DRM_OK;
    }
    NcclApi &N = nccl();
    if (mode == DRM_XCHG_FUSED) {
        rc = publish_peers(h, r);
        if (rc != DRM_OK) return rc;
        SearchDispatch disp;
        disp.world = W; disp.my_shard = r.rank; disp.slice_cap = r.slice_cap;
        for (int o = 0; o <= W; ++o) disp.slice_lo[o] = slice_lo(n, o, W);
        for (int o = 0; o < W; ++o) { disp.D_peer[o] = r.D_peer[slot][o]; disp.I_peer[o] = reinterpret_cast<long long *>(r.I_peer[slot][o]); }
        rc = launch_search_exact_ex(r.dx, q_dev, n, k, ef, nullptr, nullptr, r.counters, r.s, r.id_lo, &disp);
        if (rc != DRM_OK) return rc;
        DRM_CUDA_TRY(cudaEventRecord(r.ev[1], r.s));
        // barrier: every shard's lists for this chunk are in place before anyone merges (and, one step later, nobody
        // overwrites slot `slot` before its merge has run: the next-but-one search is behind the next barrier)
        DRM_NCCL_TRY(N.AllReduce(r.flag, r.flag, 1, ncclInt, ncclSum, r.comm, r.s));
        r.bytes_rx += (int64_t)(W - 1) * cnt * k * 12;
    } else {
        rc = launch_search_exact_ex(r.dx, q_dev, n, k, ef, r.D_loc, r.I_loc, r.counters, r.s, r.id_lo, nullptr);
        if (rc != DRM_OK) return rc;
        DRM_CUDA_TRY(cudaEventRecord(r.ev[1], r.s));
        if (mode == DRM_XCHG_ALLGATHER) {
            DRM_NCCL_TRY(N.AllGather(r.D_loc, r.D_ag, (size_t)n * k, ncclFloat, r.comm, r.s));
            DRM_NCCL_TRY(N.AllGather(r.I_loc, r.I_ag, (size_t)n * k, ncclInt64, r.comm, r.s));
            r.bytes_rx += (int64_t)(W - 1) * n * k * 12;
        } else {
            DRM_NCCL_TRY(N.GroupStart());
            for (int o = 0; o < W; ++o) {
                const int64_t olo = slice_lo(n, o, W), ocnt = slice_lo(n, o + 1, W) - olo;
                if (ocnt > 0) {
                    DRM_NCCL_TRY(N.Send(r.D_loc + (size_t)olo * k, (size_t)ocnt * k, ncclFloat, o, r.comm, r.s));
                    DRM_NCCL_TRY(N.Send(r.I_loc + (size_t)olo * k, (size_t)ocnt * k, ncclInt64, o, r.comm, r.s));
                }
                if (cnt > 0) {
                    DRM_NCCL_TRY(N.Recv(r.D_rx[slot] + (size_t)o * r.slice_cap * k, (size_t)cnt * k, ncclFloat, o, r.comm, r.s));
                    DRM_NCCL_TRY(N.Recv(r.I_rx[slot] + (size_t)o * r.slice_cap * k, (size_t)cnt * k, ncclInt64, o, r.comm, r.s));
                }
            }
            DRM_NCCL_TRY(N.GroupEnd());
            r.bytes_rx += (int64_t)(W - 1) * cnt * k * 12;
        }
    }
    DRM_CUDA_TRY(cudaEventRecord(r.ev[2], r.s));
    if (cnt > 0) {
        if (mode == DRM_XCHG_ALLGATHER)
            rc = launch_merge_topk(r.D_ag + (size_t)lo * k, r.I_ag + (size_t)lo * k, W, cnt, k, r.D_m[slot], r.I_m[slot], r.s, n);
        else
            rc = launch_merge_topk(r.D_rx[slot], r.I_rx[slot], W, cnt, k, r.D_m[slot], r.I_m[slot], r.s, r.slice_cap);
        if (rc != DRM_OK) return rc;
    }
    DRM_CUDA_TRY(cudaEventRecord(r.ev[3], r.s));
    *D_out = r.D_m[slot]; *I_out = r.I_m[slot];
    return DRM_OK;
}

int open_rank_common(ShardRank &r, const char *file, int64_t id_lo, int device, int rank, int world)
{
    HostIndex hix;
    int rc = read_index_file(file, hix);
    if (rc != DRM_OK) return rc;
    r.device = device; r.rank = rank; r.world = world; r.id_lo = id_lo;
    rc = build_device_index(hix, device, r.dx);
    if (rc != DRM_OK) return rc;
    DRM_CUDA_TRY(cudaSetDevice(device));
    DRM_CUDA_TRY(cudaStreamCreateWithFlags(&r.s, cudaStreamNonBlocking));
    DRM_CUDA_TRY(cudaStreamCreateWithFlags(&r.s_copy, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
        DRM_CUDA_TRY(cudaEventCreateWithFlags(&r.q_ready[b], cudaEventDisableTiming));
        DRM_CUDA_TRY(cudaEventCreateWithFlags(&r.q_free[b], cudaEventDisableTiming));
    }
    for (auto &set : r.evr)
        for (auto &e : set) DRM_CUDA_TRY(cudaEventCreate(&e));
    DRM_CUDA_TRY(cudaStreamCreateWithFlags(&r.s_out, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
        DRM_CUDA_TRY(cudaEventCreateWithFlags(&r.out_done[b], cudaEventDisableTiming));
        DRM_CUDA_TRY(cudaEventCreateWithFlags(&r.rel_copied[b], cudaEventDisableTiming));
        DRM_CUDA_TRY(cudaEventCreateWithFlags(&r.res_ready[b], cudaEventDisableTiming));
    }
    DRM_CUDA_TRY(cudaMalloc(&r.flag, sizeof(int)));
    DRM_CUDA_TRY(cudaMemset(r.flag, 0, sizeof(int)));
    DRM_CUDA_TRY(cudaMalloc(&r.counters, 4 * sizeof(int64_t)));
    DRM_CUDA_TRY(cudaMemset(r.counters, 0, 4 * sizeof(int64_t)));
    DRM_CUDA_TRY(cudaMalloc(&r.rmin, sizeof(int32_t)));
    return DRM_OK;
}

}  // namespace
}  // namespace drm

using namespace drm;

extern "C" {

int drm_shards_unique_id(void *id128)
{
    if (!id128) return fail(DRM_EINVAL, "null id buffer");
    NcclApi &N = nccl();
    if (!N.h) return fail(DRM_EUNSUPPORTED, N.err);
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    DRM_NCCL_TRY(N.GetUniqueId(&id));
    memcpy(id128, &id, sizeof id);
    return DRM_OK;
}

void drm_shards_close(drm_shards *h)
{
    if (!h) return;
    for (auto &up : h->local) {
        ShardRank &r = *up;
        if (r.device >= 0) { cudaSetDevice(r.device); if (r.s) cudaStreamSynchronize(r.s); if (r.s_out) cudaStreamSynchronize(r.s_out); }
    }
    for (auto &up : h->local) {
        ShardRank &r = *up;
        free_rank_buffers(r);
        if (r.comm && nccl().h) nccl().CommDestroy(r.comm);
        cudaFree(r.flag); cudaFree(r.counters); cudaFree(r.rmin);
        for (int b = 0; b < 2; ++b) {
            cudaFree(r.rq[b]); cudaFree(r.rqo[b]); cudaFree(r.rsc[b]); cudaFree(r.rid[b]);
            if (r.rel_h[b]) cudaFreeHost(r.rel_h[b]);
            if (r.q_ready[b]) cudaEventDestroy(r.q_ready[b]);
            if (r.q_free[b]) cudaEventDestroy(r.q_free[b]);
            if (r.out_done[b]) cudaEventDestroy(r.out_done[b]);
            if (r.rel_copied[b]) cudaEventDestroy(r.rel_copied[b]);
            if (r.res_ready[b]) cudaEventDestroy(r.res_ready[b]);
        }
        for (auto &set : r.evr)
            for (auto &e : set) if (e) cudaEventDestroy(e);
        if (r.s_out) cudaStreamDestroy(r.s_out);
        if (r.s) cudaStreamDestroy(r.s);
        if (r.s_copy) cudaStreamDestroy(r.s_copy);
        free_device_index(r.dx);
    }
    delete h;
}

int drm_shards_open(const char *const *index_files, const int64_t *id_lo, const int *devices, int n_shards, drm_shards **out)
{
    if (!out) return fail(DRM_EINVAL, "null output handle");
    *out = nullptr;
    if (!index_files || !id_lo || !devices || n_shards < 1 || n_shards > MAX_SHARDS) return fail(DRM_EINVAL, "bad arguments to drm_shards_open (1..16 shards)");
    for (int i = 0; i < n_shards; ++i) {  // the reference's message for a missing file, before anything touches a device
        FILE *f = index_files[i] ? fopen(index_files[i], "rb") : nullptr;
        if (!f) return fail(DRM_EIO, std::string("Index file does not exist: ") + (index_files[i] ? index_files[i] : "(null)"));
        fclose(f);
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { cudaGetLastError(); return fail(DRM_ECUDA, "no CUDA device: libdrm_cuda has no CPU fallback"); }
    for (int i = 0; i < n_shards; ++i) {
        if (devices[i] < 0 || devices[i] >= ndev) return fail(DRM_EINVAL, "device ordinal out of range");
        for (int j = 0; j < i; ++j)
            if (devices[j] == devices[i]) return fail(DRM_EINVAL, "drm_shards_open needs one distinct GPU per shard");
    }
    std::unique_ptr<drm_shards, void (*)(drm_shards *)> h(new drm_shards(), drm_shards_close);
    h->world = n_shards; h->one_process = true;
    std::vector<int> rcs((size_t)n_shards, DRM_OK);
    std::vector<std::string> errs((size_t)n_shards);
    for (int i = 0; i < n_shards; ++i) h->local.emplace_back(new ShardRank());
    {
        std::vector<std::thread> th;
        for (int i = 0; i < n_shards; ++i)
            th.emplace_back([&, i] {
                rcs[(size_t)i] = open_rank_common(*h->local[(size_t)i], index_files[i], id_lo[i], devices[i], i, n_shards);
                if (rcs[(size_t)i] != DRM_OK) errs[(size_t)i] = last_error_string();
            });
        for (auto &t : th) t.join();
    }
    for (int i = 0; i < n_shards; ++i)
        if (rcs[(size_t)i] != DRM_OK) return fail(rcs[(size_t)i], errs[(size_t)i]);
    if (n_shards > 1) {
        NcclApi &N = nccl();
        if (!N.h) return fail(DRM_EUNSUPPORTED, N.err);
        std::vector<ncclComm_t> comms((size_t)n_shards);
        DRM_NCCL_TRY(N.CommInitAll(comms.data(), n_shards, devices));
        for (int i = 0; i < n_shards; ++i) h->local[(size_t)i]->comm = comms[(size_t)i];
        bool ok = true;
        for (int i = 0; i < n_shards && ok; ++i) {
            DRM_CUDA_TRY(cudaSetDevice(devices[i]));
            for (int j = 0; j < n_shards; ++j) {
                if (i == j) continue;
                int can = 0;
                DRM_CUDA_TRY(cudaDeviceCanAccessPeer(&can, devices[i], devices[j]));
                if (!can) { ok = false; break; }
                cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else if (e != cudaSuccess) { cudaGetLastError(); ok = false; break; }
            }
        }
        h->peer_ok = ok;
    }
    *out = h.release();
    return DRM_OK;
}

int drm_shards_open_rank(const char *index_file, int64_t id_lo, int device, int rank, int world, const void *id128, drm_shards **out)
{
    if (!out) return fail(DRM_EINVAL, "null output handle");
    *out = nullptr;
    if (!index_file || world < 1 || world > MAX_SHARDS || rank < 0 || rank >= world || (world > 1 && !id128)) return fail(DRM_EINVAL, "bad arguments to drm_shards_open_rank");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { cudaGetLastError(); return fail(DRM_ECUDA, "no CUDA device: libdrm_cuda has no CPU fallback"); }
    std::unique_ptr<drm_shards, void (*)(drm_shards *)> h(new drm_shards(), drm_shards_close);
    h->world = world; h->one_process = false;
    h->local.emplace_back(new ShardRank());
    int rc = open_rank_common(*h->local[0], index_file, id_lo, device, rank, world);
    if (rc != DRM_OK) return rc;
    if (world > 1) {
        NcclApi &N = nccl();
        if (!N.h) return fail(DRM_EUNSUPPORTED, N.err);
        ncclUniqueId id;
        memcpy(&id, id128, sizeof id);
        DRM_CUDA_TRY(cudaSetDevice(device));
        DRM_NCCL_TRY(N.CommInitRank(&h->local[0]->comm, world, id, rank));
        h->peer_ok = getenv("DRM_SHARDS_NO_IPC") == nullptr;  // cudaIpc between the GPUs of one box; a failure to open falls back loudly
    }
    *out = h.release();
    return DRM_OK;
}

int drm_shards_info(const drm_shards *h, int64_t *info)
{
    if (!h || !info) return fail(DRM_EINVAL, "null argument");
    info[0] = h->world; info[1] = (int64_t)h->local.size(); info[2] = h->one_process; info[3] = h->peer_ok;
    int64_t nt = 0, bytes = 0;
    for (auto &r : h->local) { nt += r->dx.ntotal; bytes += (int64_t)r->dx.bytes; }
    info[4] = nt; info[5] = bytes; info[6] = h->local[0]->rank; info[7] = h->local[0]->id_lo;
    int v = 0;
    if (nccl().h) nccl().GetVersion(&v);
    info[8] = v;
    return DRM_OK;
}

int drm_shards_step_device(drm_shards *h, int local_rank, const float *q_dev, int64_t n, int d, int k, int ef, int exchange, void *stream_to_wait,
                           const float **D_merged, const int64_t **I_merged, int64_t *slice_lo_out, int64_t *slice_n_out, void **step_stream)
{
    if (!h || local_rank < 0 || local_rank >= (int)h->local.size() || !q_dev || !D_merged || !I_merged || !slice_lo_out || !slice_n_out)
        return fail(DRM_EINVAL, "bad arguments to drm_shards_step_device");
    ShardRank &r = *h->local[(size_t)local_rank];
    if (d != r.dx.d) return fail(DRM_EINVAL, "query dimension does not match the index");
    if (n <= 0) return fail(DRM_EEMPTY, "Query data is empty");
    (void)stream_to_wait;
    int rc = rank_step(h, r, q_dev, n, k, ef, exchange, D_merged, I_merged, slice_lo_out, slice_n_out);
    if (step_stream) *step_stream = r.s;
    return rc;
}

int drm_shards_stats(drm_shards *h, int local_rank, int reset, drm_shard_stats *st)
{
    if (!h || local_rank < 0 || local_rank >= (int)h->local.size() || !st) return fail(DRM_EINVAL, "bad arguments to drm_shards_stats");
    ShardRank &r = *h->local[(size_t)local_rank];
    DRM_CUDA_TRY(cudaSetDevice(r.device));
    DRM_CUDA_TRY(cudaStreamSynchronize(r.s));
    harvest_all(r);
    int64_t c[4] = {0, 0, 0, 0};
    DRM_CUDA_TRY(cudaMemcpy(c, r.counters, sizeof c, cudaMemcpyDeviceToHost));
    st->search_ms = r.ms_search; st->exchange_ms = r.ms_xchg; st->merge_ms = r.ms_merge; st->rerank_ms = r.ms_rerank;
    st->nhops0 = c[0]; st->nhops_up = c[1]; st->ndis_eval = c[2]; st->bytes_received = r.bytes_rx; st->steps = r.step;
    if (reset) {
        r.ms_search = r.ms_xchg = r.ms_merge = r.ms_rerank = 0; r.bytes_rx = 0;
        DRM_CUDA_TRY(cudaMemset(r.counters, 0, sizeof c));
    }
    return DRM_OK;
}

// Host-buffer entry point.  One process: W threads, one per GPU, each walking the chunks of the batch; every rank uploads
// the whole chunk of embeddings (its own PCIe link), searches, exchanges, merges its slice, optionally SW-reranks it from
// its replica of `ref`, and copies its rows of the results back.  One process per GPU: the same for the single local rank -
// only this rank's rows of D / I / sw_* are written.
int drm_shards_search_rerank(drm_shards *h, drm_ref *ref, int mode, const float *q, int64_t n, int d, int k_search, int ef, const char *q_flat,
                             const int64_t *q_off, int64_t stride, int64_t k, int64_t k_clusters, int exchange, int64_t chunk, float *D,
                             int64_t *I, int32_t *sw_scores, uint64_t *sw_ids, drm_shard_stats *stats)
{
    if (!h || !q) return fail(DRM_EINVAL, "bad arguments to drm_shards_search_rerank");
    if (n <= 0) return fail(DRM_EEMPTY, "Query data is empty");
    if (d != h->local[0]->dx.d) return fail(DRM_EINVAL, "query dimension does not match the index");
    if (exchange < 0 || exchange > 2) return fail(DRM_EINVAL, "exchange mode must be 0 (all-gather), 1 (all-to-all) or 2 (fused)");
    const bool rerank = ref != nullptr && sw_scores != nullptr && sw_ids != nullptr;
    if (rerank && (!q_flat || !q_off)) return fail(DRM_EINVAL, "rerank needs the read sequences");
    if (rerank && k > k_clusters * 2 * stride) return fail(DRM_EKTOOLARGE, "Final k too large. Ensure k < k_clusters * 2 * stride to have enough candidates.");
    if (chunk <= 0) chunk = 262144;
    if (chunk > n) chunk = n;
    std::lock_guard<std::mutex> lk(h->mu);
    const int L = (int)h->local.size();
    const int eff_mode = (exchange == DRM_XCHG_FUSED && h->world > 1 && !h->peer_ok) ? DRM_XCHG_ALLTOALL : exchange;
    // buffers of every local rank first, then the peer tables: a rank must not read another rank's pointers before they exist
    for (int li = 0; li < L; ++li) {
        ShardRank &r = *h->local[(size_t)li];
        int rc = ensure_buffers(h, r, std::max<int64_t>(chunk, r.cap_chunk), k_search, d, eff_mode);
        if (rc != DRM_OK) return rc;
    }
    if (eff_mode == DRM_XCHG_FUSED && h->world > 1)
        for (int li = 0; li < L; ++li) {
            int rc = publish_peers(h, *h->local[(size_t)li]);
            if (rc != DRM_OK) return rc;
        }
    std::vector<int> rcs((size_t)L, DRM_OK);
    std::vector<std::string> errs((size_t)L);
    std::vector<int32_t> minc((size_t)L, 0x7fffffff);
    auto work = [&](int li) -> int {
        ShardRank &r = *h->local[(size_t)li];
        const int W = r.world;
        DRM_CUDA_TRY(cudaSetDevice(r.device));
        const DeviceRef *dref = nullptr;
        if (rerank) {
            for (auto &dr : ref->rep) if (dr.device == r.device) dref = &dr;
            if (!dref) return fail(DRM_EINVAL, "the reference handle has no replica on a shard's device");
        }
        int rc = DRM_OK;
        (void)W;
        const int32_t big = 0x7fffffff;
        DRM_CUDA_TRY(cudaMemcpyAsync(r.rmin, &big, sizeof big, cudaMemcpyHostToDevice, r.s));
        if (rerank) {  // staging for the largest slice of this call, two sets
            size_t need_rq = 0, need_rows = 0;
            for (int64_t c0 = 0; c0 < n; c0 += chunk) {
                const int64_t nc = std::min(chunk, n - c0);
                const int64_t lo = slice_lo(nc, r.rank, r.world), cnt = slice_lo(nc, r.rank + 1, r.world) - lo;
                need_rows = std::max(need_rows, (size_t)cnt);
                need_rq = std::max(need_rq, (size_t)(q_off[c0 + lo + cnt] - q_off[c0 + lo]));
            }
            if (need_rq + 16 > r.cap_rq || need_rows + 1 > r.cap_rqo || need_rows * (size_t)k > r.cap_rout) {
                DRM_CUDA_TRY(cudaStreamSynchronize(r.s));
                DRM_CUDA_TRY(cudaStreamSynchronize(r.s_out));
                r.cap_rq = std::max(r.cap_rq, need_rq + 4096); r.cap_rqo = std::max(r.cap_rqo, need_rows + 64); r.cap_rout = std::max(r.cap_rout, need_rows * (size_t)k + 64);
                for (int b = 0; b < 2; ++b) {
                    cudaFree(r.rq[b]); cudaFree(r.rqo[b]); cudaFree(r.rsc[b]); cudaFree(r.rid[b]);
                    if (r.rel_h[b]) cudaFreeHost(r.rel_h[b]);
                    r.rq[b] = nullptr; r.rqo[b] = nullptr; r.rsc[b] = nullptr; r.rid[b] = nullptr; r.rel_h[b] = nullptr;
                    DRM_CUDA_TRY(cudaMalloc(&r.rq[b], r.cap_rq));
                    DRM_CUDA_TRY(cudaMalloc(&r.rqo[b], r.cap_rqo * sizeof(int64_t)));
                    DRM_CUDA_TRY(cudaMalloc(&r.rsc[b], r.cap_rout * sizeof(int32_t)));
                    DRM_CUDA_TRY(cudaMalloc(&r.rid[b], r.cap_rout * sizeof(uint64_t)));
                    DRM_CUDA_TRY(cudaHostAlloc(&r.rel_h[b], r.cap_rqo * sizeof(int64_t), cudaHostAllocDefault));
                }
            }
        }
        int64_t c_idx = 0;
        for (int64_t c0 = 0; c0 < n; c0 += chunk, ++c_idx) {
            const int64_t nc = std::min(chunk, n - c0);
            const int qb = (int)(c_idx & 1);
            // upload of the chunk's embeddings on the copy stream, ordered after the search that last read this buffer
            DRM_CUDA_TRY(cudaStreamWaitEvent(r.s_copy, r.q_free[qb], 0));
            DRM_CUDA_TRY(cudaMemcpyAsync(r.q[qb], q + (size_t)c0 * d, (size_t)nc * d * sizeof(float), cudaMemcpyHostToDevice, r.s_copy));
            DRM_CUDA_TRY(cudaEventRecord(r.q_ready[qb], r.s_copy));
            DRM_CUDA_TRY(cudaStreamWaitEvent(r.s, r.q_ready[qb], 0));
            // the merge buffers and the rerank outputs this step will overwrite: their copies home (two steps ago) must be done
            const int ob = (int)(r.step & 1);
            DRM_CUDA_TRY(cudaStreamWaitEvent(r.s, r.out_done[ob], 0));
            const float *Dm = nullptr; const int64_t *Im = nullptr;
            int64_t lo = 0, cnt = 0;
            rc = rank_step(h, r, r.q[qb], nc, k_search, ef, exchange, &Dm, &Im, &lo, &cnt);
            if (rc != DRM_OK) return rc;
            DRM_CUDA_TRY(cudaEventRecord(r.q_free[qb], r.s));
            const int64_t g0 = c0 + lo;
            if (cnt > 0 && rerank) {
                const int64_t b0 = q_off[g0], b1 = q_off[g0 + cnt];
                DRM_CUDA_TRY(cudaEventSynchronize(r.rel_copied[ob]));  // the host staging of two steps ago has been read
                int64_t *rel = r.rel_h[ob];
                int maxlen = 0;
                for (int64_t i = 0; i <= cnt; ++i) rel[i] = q_off[g0 + i] - b0;
                for (int64_t i = 0; i < cnt; ++i) maxlen = std::max<int>(maxlen, (int)(rel[i + 1] - rel[i]));
                DRM_CUDA_TRY(cudaMemcpyAsync(r.rq[ob], q_flat + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, r.s));
                DRM_CUDA_TRY(cudaMemcpyAsync(r.rqo[ob], rel, (size_t)(cnt + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, r.s));
                DRM_CUDA_TRY(cudaEventRecord(r.rel_copied[ob], r.s));
                RerankArgs a{mode, Im, cnt, (int64_t)k_search, r.rq[ob], r.rqo[ob], maxlen, stride, k, k_clusters, r.rsc[ob], r.rid[ob], nullptr, r.rmin, nullptr};
                rc = launch_rerank(*dref, a, r.s);
                if (rc != DRM_OK) return rc;
                DRM_CUDA_TRY(cudaEventRecord(r.ev[4], r.s));
                r.ev_last[r.ev_cur] = 4;
            }
            // results go home on their own stream while the next chunk is searched
            DRM_CUDA_TRY(cudaEventRecord(r.res_ready[ob], r.s));
            DRM_CUDA_TRY(cudaStreamWaitEvent(r.s_out, r.res_ready[ob], 0));
            if (cnt > 0 && rerank) {
                DRM_CUDA_TRY(cudaMemcpyAsync(sw_scores + (size_t)g0 * k, r.rsc[ob], (size_t)cnt * k * sizeof(int32_t), cudaMemcpyDeviceToHost, r.s_out));
                DRM_CUDA_TRY(cudaMemcpyAsync(sw_ids + (size_t)g0 * k, r.rid[ob], (size_t)cnt * k * sizeof(uint64_t), cudaMemcpyDeviceToHost, r.s_out));
            }
            if (cnt > 0 && D) DRM_CUDA_TRY(cudaMemcpyAsync(D + (size_t)g0 * k_search, Dm, (size_t)cnt * k_search * sizeof(float), cudaMemcpyDeviceToHost, r.s_out));
            if (cnt > 0 && I) DRM_CUDA_TRY(cudaMemcpyAsync(I + (size_t)g0 * k_search, Im, (size_t)cnt * k_search * sizeof(int64_t), cudaMemcpyDeviceToHost, r.s_out));
            DRM_CUDA_TRY(cudaEventRecord(r.out_done[ob], r.s_out));
        }
        DRM_CUDA_TRY(cudaStreamSynchronize(r.s));
        DRM_CUDA_TRY(cudaStreamSynchronize(r.s_out));
        harvest_all(r);
        if (rerank) {
            int32_t m = 0x7fffffff;
            DRM_CUDA_TRY(cudaMemcpy(&m, r.rmin, sizeof m, cudaMemcpyDeviceToHost));
            minc[(size_t)li] = m;
        }
        return DRM_OK;
    };
    if (L == 1) {
        rcs[0] = work(0);
        if (rcs[0] != DRM_OK) errs[0] = last_error_string();
    } else {
        std::vector<std::thread> th;
        for (int li = 0; li < L; ++li)
            th.emplace_back([&, li] { rcs[(size_t)li] = work(li); if (rcs[(size_t)li] != DRM_OK) errs[(size_t)li] = last_error_string(); });
        for (auto &t : th) t.join();
    }
    for (int li = 0; li < L; ++li)
        if (rcs[(size_t)li] != DRM_OK) return fail(rcs[(size_t)li], errs[(size_t)li]);
    if (rerank) {
        int32_t m = 0x7fffffff;
        for (int32_t v : minc) m = std::min(m, v);
        if ((int64_t)m < k) return fail(DRM_ENOTENOUGH, "Not enough candidates (" + std::to_string(m) + " < " + std::to_string(k) + ")");
    }
    if (stats) return drm_shards_stats(h, 0, 0, stats);
    return DRM_OK;
}

}  // extern "C"
