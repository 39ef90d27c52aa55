// Shared declarations of libdrm_cuda.so (internal; the public boundary is include/drm_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/drm_b200.h"

namespace drm {

void set_error(const std::string &msg);
int fail(int code, const std::string &msg);
const std::string &last_error_string();

#define DRM_CUDA_TRY(expr)                                                                              \
    do {                                                                                                \
        cudaError_t e__ = (expr);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            return ::drm::fail(DRM_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));         \
    } while (0)

// ---------------------------------------------------------------------------------------------
// Host image of a FAISS IndexHNSWPQ ("IHNp") file — the arrays exactly as FAISS serialises them
// (layout: SURVEY.md §8c item 1; reference call site src/main.cpp:236-237).
struct HostIndex {
    int32_t d = 0;
    int64_t ntotal = 0;
    int32_t metric = 1;
    std::vector<double> assign_probas;
    std::vector<int32_t> cum;        // cum_nneighbor_per_level
    std::vector<int32_t> levels;     // top level + 1 per vertex
    std::vector<uint64_t> offsets;   // ntotal + 1
    std::vector<int32_t> neighbors;  // -1 padded per level slot
    int32_t entry_point = -1, max_level = -1, efConstruction = 40, efSearch = 16, upper_beam = 1;
    uint64_t M = 0, nbits = 0, ksub = 0, dsub = 0, code_size = 0;
    std::vector<float> centroids;  // [M][ksub][dsub]
    std::vector<uint8_t> codes;    // ntotal x code_size
    int32_t search_type = 0;       // PQ search_type (ST_PQ)
    uint8_t encode_signs = 0;
    int32_t polysemous_ht = 0;
    int M_hnsw() const { return cum.size() > 1 ? cum[1] / 2 : 0; }
};

int read_index_file(const char *path, HostIndex &ix);         // DRM_OK or error (message set)
int write_index_file(const char *path, const HostIndex &ix);  // same layout, used by the builder

// Pieces of the index builder shared by the host (builder.cpp) and GPU (build_gpu.cu) builders, so that both draw the
// same training sample, k-means start, vertex levels and insertion order from a seed.
void builder_level_law(HostIndex &ix, int64_t n, int M_hnsw, uint32_t seed);
void builder_insertion_order(const HostIndex &ix, uint32_t seed, std::vector<int64_t> &order);
void builder_training_pick(int64_t n_train, uint32_t seed, std::vector<int64_t> &pick);
void builder_kmeans_init_rows(int64_t n, int k, uint32_t seed, std::vector<int64_t> &rows);
void builder_kmeans_update(const double *sum, int64_t *cnt, int k, int dsub, float *cent);
int build_index_cpu(const float *x, int64_t n, int d, const float *train, int64_t n_train, int M_pq, int nbits, int M_hnsw,
                    int efc, int threads, uint32_t seed, HostIndex &ix);

// ---------------------------------------------------------------------------------------------
// Device image of the index, laid out for one-round-trip hops (DESIGN.md "data layout in HBM"):
//   rows0 : ntotal rows; row v = [W0 x int32 level-0 neighbour ids (-1 padded)]
//                                [W0 x code_size bytes: the PQ code of each of those neighbours]
//   up    : one row per (vertex, level>=1), the levels of a vertex contiguous and ascending;
//           row = [int32 vertex id, int32 level, 8 pad bytes][Wu x int32 neighbour ROW indices at the
//                  same level (-1 padded)][Wu x code_size bytes of neighbour codes]
struct DeviceIndex {
    int device = -1;
    int d = 0, M = 0, ksub = 0, dsub = 0, code_size = 0;
    int W0 = 0, Wu = 0;        // row widths: 2*M_hnsw and M_hnsw
    int row0_bytes = 0, rowu_bytes = 0;
    int code0_off = 0, codeu_off = 0;  // byte offset of the code block inside a row (16-byte aligned)
    int64_t ntotal = 0;
    int max_level = 0;
    int64_t entry_vertex = -1;
    int64_t entry_row = -1;    // upper row of the entry point at max_level; -1 when max_level == 0
    uint8_t entry_code[64] = {0};
    float *centroids = nullptr;
    float *centroids_t = nullptr;  // transposed copy [M][dsub/4][ksub][4] for coalesced LUT builds (dsub % 4 == 0)
    uint8_t *rows0 = nullptr;
    uint8_t *up = nullptr;
    int64_t n_up_rows = 0;
    int64_t *counters = nullptr;  // 3 x int64 scratch for host-pointer calls
    size_t bytes = 0;
};

int build_device_index(const HostIndex &h, int device, DeviceIndex &out);
void free_device_index(DeviceIndex &dx);

// Kernel launchers (search_kernel.cu)
int launch_search(const DeviceIndex &dx, const float *q_dev, int64_t n, int k, int ef, float *D_dev, int64_t *I_dev,
                  int64_t *counters_dev, cudaStream_t stream);
// FAISS-exact kernel (search_exact.cu): same contract; launch_search routes to it when the shape is supported
bool search_exact_supports(const DeviceIndex &dx, int k, int ef);
int launch_search_exact(const DeviceIndex &dx, const float *q_dev, int64_t n, int k, int ef, float *D_dev, int64_t *I_dev,
                        int64_t *counters_dev, cudaStream_t stream);
// Position-sharded search (shards.cu): labels leave the kernel already global (+ id_add), and with a dispatch table the
// kernel's epilogue writes every read's list straight into the buffers of the rank that merges that read - local memory
// or a peer GPU's over NVLink - so the exchange step of the sharded path is fused into the search.
constexpr int MAX_SHARDS = 16;
struct SearchDispatch {
    int world = 0, my_shard = 0;
    long long slice_cap = 0;              // rows of one (shard, owner) block in the owner's buffers
    long long slice_lo[MAX_SHARDS + 1];   // owner o merges reads [slice_lo[o], slice_lo[o + 1]) of this launch
    float *D_peer[MAX_SHARDS];            // owner o's list buffers [world][slice_cap][k] (this shard writes block my_shard)
    long long *I_peer[MAX_SHARDS];
};
int launch_search_exact_ex(const DeviceIndex &dx, const float *q_dev, int64_t n, int k, int ef, float *D_dev, int64_t *I_dev,
                           int64_t *counters_dev, cudaStream_t stream, int64_t id_add, const SearchDispatch *disp);

// ---------------------------------------------------------------------------------------------
// Reference windows on the device (sw_kernel.cu)
struct DeviceRef {
    int device = -1;
    int form = 0;            // 0 = genome, 1 = window table
    int ref_len = 0;
    int64_t len = 0;         // genome bases, or number of windows
    uint8_t *data = nullptr; // base classes 0..4 (A,C,G,T,N), 7 = any other byte (table form keeps raw too)
    uint8_t *data_alloc = nullptr; // the allocation behind `data`: REF_PAD bytes of class 4 either side, so the rerank kernel's
                                   // aligned word reads around a window never leave it
    uint8_t *raw = nullptr;  // raw bytes (only kept when has_other: the generic byte-equality path)
    bool has_other = false;  // some byte outside ACGTN -> generic kernel
};

constexpr int REF_PAD = 32;

struct RerankArgs {
    int mode;  // 0 static bound rule, 1 dynamic
    const int64_t *neighbors; int64_t n, k_in;
    const char *q_flat; const int64_t *q_off; int max_qlen;
    int64_t stride, k, k_clusters;
    int32_t *out_scores; uint64_t *out_ids; int32_t *n_cands; int32_t *min_cands; int64_t *cells;
};
int launch_rerank(const DeviceRef &r, const RerankArgs &a, cudaStream_t stream);
int launch_sw_pairs(const char *a_flat, const int64_t *a_off, const char *b_flat, const int64_t *b_off, int64_t n,
                    int max_len_b, int32_t *scores, cudaStream_t stream);
int launch_merge_topk(const float *D_all, const int64_t *I_all, int n_shards, int64_t n, int k, float *D_out,
                      int64_t *I_out, cudaStream_t stream, int64_t shard_stride = -1 /* rows between two shards' blocks; -1 = n */);
int launch_classify(const uint8_t *raw, uint8_t *cls, int64_t n, int *has_other_dev, cudaStream_t stream);

// SW with traceback (align_kernel.cu); every pointer is a device pointer.
int launch_sw_align(const DeviceRef &ref, const unsigned long long *ids, int64_t n_reads, int per_read, const char *q_flat,
                    const int64_t *q_off, int max_qlen, int strip, int max_ops, int *score, int *ref_beg, int *ref_end, int *q_beg,
                    int *q_end, int *n_ops, int *nm, unsigned *ops, cudaStream_t stream);

// Input side on the device (input_kernels.cu).  Output buffers are fresh cudaMalloc allocations on `device` owned by the caller.
int clean_fasta_device(const char *data, int64_t size, int device, uint8_t **d_genome, int64_t *len);
int parse_fastq_device(const char *data, int64_t size, int device, char **d_flat, int64_t **d_off, int64_t *n_reads,
                       int64_t *flat_bytes, int *max_len);

// ---------------------------------------------------------------------------------------------
// Embedder: 3-mer tokeniser + token embedding + 2-layer bidirectional GRU (hidden 64) -> 128-d,
// the network in the reference's models/*.xml that src/inference/vectorize.cpp:34-141 runs through OpenVINO.
// ir_model.cpp reads the IR; embed_kernel.cu holds the device side.
struct EmbedWeightsHost {  // fp32, gate order z, r, h as GRUSequence stores them
    int vocab = 0;
    std::vector<float> emb;         // [vocab][64]
    std::vector<float> W1, R1, B1;  // [2][192][64], [2][192][64], [2][256]
    std::vector<float> W2, R2, B2;  // [2][192][128], [2][192][64], [2][256]
};
int read_ir_model(const char *xml_path, const char *bin_path, EmbedWeightsHost &w);
float half_bits_to_float(uint16_t h);

struct DeviceEmbedder {
    int device = -1;
    int max_len = 123;  // Config::Inference::MAX_LEN (includes/utils/config.hpp:21)
    float *P1 = nullptr;   // [2][97][192] layer-1 input projection of every token the tokeniser can emit (+ bias)
    float *R1 = nullptr, *bR1 = nullptr;  // [2][192][64], [2][64]
    float *W2 = nullptr, *b2 = nullptr;   // [2][192][128], [2][192]
    float *R2 = nullptr, *bR2 = nullptr;  // [2][192][64], [2][64]
    int64_t chunk = 0;     // sequences per pass; the scratch below is sized for it
    float *Y1 = nullptr;   // [chunk][max_len][128] layer-1 states, forward | backward
    float *X2 = nullptr;   // [chunk][max_len][2][192] layer-2 input projections
    float *out = nullptr;  // [chunk][128] staging for host-pointer calls
    // tensor-core form (embed_kernel.cu, gru_tc_kernel): fp16 operand images of the recurrent weights
    uint16_t *Bw1 = nullptr;    // [2][16][192][8]  [W1 | R1] per direction, [k-group][row][8 halves]
    uint16_t *Bw2 = nullptr;    // [2][24][192][8]  [W2 | R2] per direction
    uint16_t *emb16 = nullptr;  // [97][64]         the embedding rows the tokeniser can select
    float *bias1 = nullptr;     // [2][192 + 64]    input biases | recurrent bias of the candidate gate, per direction
    float *bias2 = nullptr;     // [2][192 + 64]
    uint16_t *Y1h = nullptr;    // [chunk_tc / 128][max_len][hi | lo][16 k-groups][128 rows][8 halves]: layer-1 states as the next layer's operand images
    int64_t chunk_tc = 0;
    bool tc_exact = false;      // every recurrent weight is an fp16 number (true for the reference's IR): the tensor-core form is exact in B
};
struct EmbedSource {
    int mode = 0;  // 0: byte strings flat/off; 1: reference windows by window id (find_sequence ids)
    const char *flat = nullptr; const int64_t *off = nullptr;
    const uint8_t *cls = nullptr; int form = 0; int64_t ref_n = 0; int ref_len = 0;  // DeviceRef::data
    unsigned long long first_id = 0, id_step = 1;   // window i has id first_id + i * id_step ...
    const unsigned long long *ids = nullptr;        // ... unless an id list (device) is given
    int decorate = 1;  // 1: '<' + window + '>' (the strings format_fasta builds); 0: the bare window find_sequence returns
};
int embedder_upload(const EmbedWeightsHost &w, int device, int max_len, DeviceEmbedder &e);
void embedder_free(DeviceEmbedder &e);
// Embeds sequences [first, first + n) of `src` into out_dev (n x 128 floats on the embedder's device).
void tokenize_host(const char *flat, const int64_t *off, int64_t n, int T, uint16_t *tokens);
int launch_embed(DeviceEmbedder &e, const EmbedSource &src, int64_t first, int64_t n, float *out_dev, cudaStream_t stream);

// ---------------------------------------------------------------------------------------------
// L2 rerank stage (post_process_l2_static / _dynamic + batch_reranker; l2_plan.cpp, l2_kernel.cu)
struct L2Plan {
    int64_t n = 0, k_out = 0;
    bool passthrough = false;            // dynamic variant on a dense index: search order and distances pass through
    int64_t take = 0;                    // ... the first `take` neighbours of every query
    std::vector<uint64_t> unique;        // window ids to embed (dynamic: find_sequence ids; static: rows of ref_seqs)
    std::vector<uint32_t> map;           // flat candidate list -> row of `unique`
    std::vector<int64_t> starts;         // n + 1 slice boundaries into the flat list, as the reference computes them
    bool direct_ids = false;             // asked for by the caller (all window ids below 2^32, embeddings of every window at hand):
                                         // `map` holds the window ids themselves, `unique` stays empty - no de-duplication pass
};
// mode 0 = static (bound = number of ref_seqs), 1 = dynamic (bound = genome length).  DRM_OK or an error with the
// reference's message.
int l2_plan(int mode, const int64_t *neighbors, int64_t n, int64_t k_in, int64_t stride, int64_t k, int64_t k_clusters,
            uint64_t bound, L2Plan &plan, bool direct_ids = false);
// libstdc++ std::partial_sort(idx, idx + k, idx + n, dist[a] < dist[b]) on the host (the code the kernel runs)
void l2_select_host(const float *dist, int64_t n, int64_t k, int64_t *idx);
// emb: unique windows' embeddings [U][128]; map/starts as in L2Plan (device copies); q: [n][128]; dist_scr / idx_scr: one
// entry per flat candidate.  Writes k distances + rows of `unique` per query (n x k) and counts[q] = k, 0 (no
// candidates) or -(number of candidates) - 1 when there are fewer than k.
int launch_l2_rerank(const float *emb, const uint32_t *map, const int64_t *starts, const float *q, int64_t n, int k,
                     float *dist_scr, int64_t *idx_scr, float *out_dist, uint32_t *out_rows, int64_t *counts,
                     cudaStream_t stream);

}  // namespace drm

// ---------------------------------------------------------------------------------------------
// The opaque handles of include/drm_b200.h (global namespace, C linkage names), shared by api.cu and shards.cu.
// Per-replica workspace of drm_search_rerank, kept in the index handle so that repeated calls reuse their streams,
// device buffers and pinned offset staging instead of paying cudaMalloc / cudaFree on every call.
struct StageSlot {
    cudaStream_t s = nullptr;
    void *q = nullptr, *D = nullptr, *I = nullptr, *qs = nullptr, *qo = nullptr, *sc = nullptr, *id = nullptr;
    int64_t *off = nullptr;  // pinned host: chunk-relative read offsets
};
struct StageWorkspace {
    static constexpr int NSLOT = 3;
    StageSlot slot[NSLOT];
    void *cnt = nullptr, *dmin = nullptr;
    size_t cap_q = 0, cap_D = 0, cap_I = 0, cap_qs = 0, cap_qo = 0, cap_sc = 0, cap_id = 0;
    std::mutex mu;
};

struct drm_index {
    std::vector<drm::DeviceIndex> rep;
    std::vector<std::unique_ptr<StageWorkspace>> ws;  // one per replica, created on first use
    int64_t info[12];
};

struct drm_ref {
    std::vector<drm::DeviceRef> rep;
    uint64_t uid = 0;  // process-unique, never reused: what caches of derived data key on (an address can be reused after a close)
};

// FASTQ reads parsed on the device (input_kernels.cu): one copy per replica.
struct ReadsRep {
    int device = -1;
    char *flat = nullptr;     // '<' + sequence + '>' of every read, back to back
    int64_t *off = nullptr;   // n + 1 offsets into flat
};
struct drm_reads {
    std::vector<ReadsRep> rep;
    int64_t n = 0, bytes = 0;
    int max_len = 0;
};

struct drm_embedder {
    drm::DeviceEmbedder dev;
    // drm_l2_rerank: embeddings of EVERY window id of one reference ([n][128] floats on dev.device), built the first time a call
    // references a sizeable share of them and reused by later calls on the same reference (keyed on drm_ref::uid and the window form)
    struct WindowCache {
        uint64_t ref_uid = 0;
        int decorate = -1;
        uint64_t n = 0;
        float *emb = nullptr;
    } wcache;
};

