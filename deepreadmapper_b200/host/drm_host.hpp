// Host side of the drop-in: the reference's C++ call signatures for the search-time hot path, implemented over
// the C ABI of libdrm_cuda.so (include/drm_b200.h).  A `pipeline` built against this header keeps the structure of
// the reference's src/main.cpp: read_index -> faiss_search -> post_process_sw_* -> save_results.
//
// Signatures mirrored (paths relative to the reference repository):
//   faiss::read_index / faiss::IndexHNSWPQ                src/main.cpp:236-237,299
//   faiss_search                                          includes/hnswpq/search.hpp:18-22, src/hnswpq/search.cpp:6-57
//   post_process_sw_static / post_process_sw_dynamic      includes/utils/post_processor.hpp:79-103
//   sw_reranker                                           includes/utils/reranker.hpp:19-23
//   calc_sw_score                                         includes/utils/metrics.hpp:22
//   Vectorizer                                            includes/inference/vectorize.hpp:14-56, src/inference/vectorize.cpp:4-141
//   load_config / save_config / save_results / read_file  includes/utils/utils.hpp:53-115
//   reverse_complement / extract_FASTA_sequence / preprocess_fasta / preprocess_fastq   includes/utils/parse_inputs.hpp
// Errors are std::runtime_error with the reference's messages; main catches once and prints "Error: <what>".
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>
#include <tuple>
#include <unordered_map>
#include <utility>
#include <variant>
#include <vector>

struct drm_index;
struct drm_ref;
struct drm_embedder;
struct drm_shards;

// includes/utils/config.hpp:6-57 (only the constants the path uses)
namespace Config {
namespace Build { constexpr int DIM = 128; constexpr int M_PQ = 8; constexpr int NBITS = 8; constexpr int M = 16; constexpr int EFC = 200;
                  constexpr float SAMPLE_RATE = 0.5f; }
namespace Inference { constexpr const char MODEL_PATH[] = "models/finetuned_sgn33-new-a-Apr6.xml"; constexpr size_t BATCH_SIZE = 100;
                      constexpr size_t MAX_LEN = 123; constexpr size_t MODEL_OUT_SIZE = 128; }
namespace Search { constexpr int EF = 128; constexpr int K = 128; constexpr int K_CLUSTERS = 5; }
}  // namespace Config

namespace faiss {
using idx_t = int64_t;
// Handle on an IHNp index resident on the GPU(s).  `delete` releases the device memory, as src/main.cpp:299 expects.
struct IndexHNSWPQ {
    drm_index *handle = nullptr;
    int d = 0;
    int64_t ntotal = 0;
    struct { int efSearch = 16; int efConstruction = 40; } hnsw;
    ~IndexHNSWPQ();
};
// Devices: every visible GPU unless DRM_DEVICES="0,1,.." restricts them (index replicated, reads sharded).
IndexHNSWPQ *read_index(const char *index_file);
}  // namespace faiss

std::pair<std::vector<std::vector<faiss::idx_t>>, std::vector<std::vector<float>>> faiss_search(
    faiss::IndexHNSWPQ *index, const std::vector<std::vector<float>> &query_data, int k = Config::Search::K, int ef = Config::Search::EF);

// Position-sharded index (additive; BASELINE.json configs[4]): one IHNp file per GPU, written by `hnswpq_index` under
// DRM_SHARDS=<W> as <prefix>/<basename>.shard<g>.index with the config keys n_shards and shard_<g>_id_lo.  One call where the
// reference calls faiss_search once (src/main.cpp:278): every shard searches every read, the per-shard lists are merged on
// the GPU that owns the read, and - when `genome` is not empty - that GPU SW-reranks its reads.  exchange = DRM_XCHG from
// the environment ("allgather" | "alltoall" | "fused", default fused).
struct ShardedIndex {
    drm_shards *handle = nullptr;
    int world = 0;
    std::vector<int> devices;
    ~ShardedIndex();
};
ShardedIndex *read_index_sharded(const std::string &index_prefix, const std::unordered_map<std::string, std::variant<size_t, float, std::string>> &config);
void sharded_search_rerank_flat(ShardedIndex *index, const float *q, int64_t n, int d, int k_search, int ef, int mode, const std::string &genome,
                                size_t ref_len, const std::vector<std::string> &query_seqs, size_t stride, size_t k, size_t k_clusters,
                                std::vector<faiss::idx_t> &I, std::vector<float> &D, std::vector<int> &sw_scores, std::vector<size_t> &sw_ids);

// Flat form used by the pipeline for large batches (no nested vectors): q is n x d row-major.
void faiss_search_flat(faiss::IndexHNSWPQ *index, const float *q, int64_t n, int d, int k, int ef, std::vector<faiss::idx_t> &I, std::vector<float> &D);

// Search + SW rerank with the FASTQ reads and the FASTA genome parsed on the device (additive; `pipeline` uses it for
// FASTQ queries unless DRM_INPUT=host).  Returns the number of reads; throws the reference's messages on mismatch.
int64_t stage_device_inputs(faiss::IndexHNSWPQ *index, const std::string &fastq_file, const std::string &fasta_file, const float *emb,
                            int64_t n_emb, int d, int k_search, int ef, int mode, size_t ref_len, size_t stride, size_t k, size_t k_clusters,
                            std::vector<faiss::idx_t> &I, std::vector<float> &D, std::vector<int> &sw_scores, std::vector<size_t> &sw_ids);

// write_sam (src/utils/utils.cpp:336-398) with real POS / CIGAR / NM (traceback kernel) and a score-gap MAPQ; the first n_out
// reranked windows of every read are written (query_seqs carry the '<' '>' decoration like the reference's).
void write_sam_aligned(const std::string &sam_file, const std::string &ref_name, const std::string &genome, size_t ref_len,
                       const std::vector<std::string> &query_seqs, const std::vector<std::string> &query_ids, const std::vector<int> &sw_scores,
                       const std::vector<size_t> &sw_ids, size_t k, size_t n_out);

// The embedder: tokeniser + network on the GPU (csrc/embed_kernel.cu).  model_path = the OpenVINO IR .xml the reference
// loads (the .bin beside it); batch_size is accepted for compatibility (the device batches by itself).
class Vectorizer {
public:
    Vectorizer(const std::string &model_path = Config::Inference::MODEL_PATH, size_t batch_size = Config::Inference::BATCH_SIZE,
               size_t max_len = Config::Inference::MAX_LEN, size_t model_out_size = Config::Inference::MODEL_OUT_SIZE);
    ~Vectorizer();
    Vectorizer(const Vectorizer &) = delete;
    Vectorizer &operator=(const Vectorizer &) = delete;
    std::vector<std::vector<float>> vectorize(const std::vector<std::string> &input, bool verbose = false);
    // flat form for large batches: out = input.size() x 128 row-major
    void vectorize_flat(const std::vector<std::string> &input, std::vector<float> &out);
    drm_embedder *handle() const { return handle_; }
private:
    drm_embedder *handle_ = nullptr;
    size_t model_out_size_;
};
// DRM_MODEL if set, else Config::Inference::MODEL_PATH (relative to the working directory, like the reference).
std::string embedder_model_path();

int calc_sw_score(const std::string &seq1, const std::string &seq2);

std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> sw_reranker(
    const std::vector<std::string> &cand_seqs, const std::vector<size_t> &cand_ids, const std::string &query_seq, size_t k);

template <typename NeighborType>
std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> post_process_sw_dynamic(
    const std::vector<std::vector<NeighborType>> &neighbors, const std::vector<std::vector<float>> &distances, const std::string &ref_genome,
    const std::vector<std::string> &query_seqs, size_t ref_len, size_t stride, size_t k, size_t k_clusters = 5);

template <typename NeighborType>
std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> post_process_sw_static(
    const std::vector<std::vector<NeighborType>> &neighbors, const std::vector<std::vector<float>> &distances,
    const std::vector<std::string> &ref_seqs, const std::vector<std::string> &query_seqs, size_t ref_len, size_t stride, size_t k,
    size_t k_clusters = 5);

// The L2 rerank variants (includes/utils/post_processor.hpp:118-141,170-181) over drm_l2_rerank: candidate windows are
// embedded on the device and ranked by embedding distance.  Every query must carry the same number of neighbours.
// (Device path awaiting its first hardware run - DESIGN.md section 4.10.)
template <typename NeighborType>
std::tuple<std::vector<std::string>, std::vector<float>, std::vector<size_t>> post_process_l2_dynamic(
    const std::vector<std::vector<NeighborType>> &neighbors, const std::vector<std::vector<float>> &distances, const std::string &ref_genome,
    const std::vector<std::string> &query_seqs, size_t ref_len, size_t stride, size_t k, const std::vector<std::vector<float>> &query_embeddings,
    Vectorizer &vectorizer, size_t k_clusters = 5);

template <typename NeighborType>
std::tuple<std::vector<std::string>, std::vector<float>, std::vector<size_t>> post_process_l2_static(
    const std::vector<std::vector<NeighborType>> &neighbors, const std::vector<std::vector<float>> &distances,
    const std::vector<std::string> &ref_seqs, const std::vector<std::string> &query_seqs, size_t ref_len, size_t stride, size_t k,
    const std::vector<std::vector<float>> &query_embeddings, Vectorizer &vectorizer, size_t k_clusters = 5);

// Flat rerank for the pipeline: neighbours n x k_in (int64), returns scores / ids n x k; candidate strings are not built.
// mode 0 = static bound rule, 1 = dynamic.  `genome` is the cleaned reference (extract_FASTA_sequence).
void post_process_sw_flat(int mode, const std::string &genome, size_t ref_len, const std::vector<faiss::idx_t> &I, int64_t n, int64_t k_in,
                          const std::vector<std::string> &query_seqs, size_t stride, size_t k, size_t k_clusters, std::vector<int> &scores,
                          std::vector<size_t> &ids);

// Flat L2 rerank for the pipeline (DRM_RERANK=l2): mode 1 = post_process_l2_dynamic over `genome`, mode 0 =
// post_process_l2_static over `ref_windows` (bare lookup-mode windows, or '<' window '>' strings).  dists / ids are flat, like the reference's final vectors.
void post_process_l2_flat(int mode, const std::string &genome, const std::vector<std::string> &ref_windows, size_t ref_len,
                          const std::vector<faiss::idx_t> &I, const std::vector<float> &D, int64_t n, int64_t k_in, const float *q_emb,
                          Vectorizer &vectorizer, size_t stride, size_t k, size_t k_clusters, std::vector<float> &dists, std::vector<size_t> &ids);

// ---- parse_inputs
std::string reverse_complement(const std::string &seq);
std::string extract_FASTA_sequence(const std::string &fasta_file);
std::pair<std::vector<std::string>, std::vector<size_t>> preprocess_fasta(const std::string &fasta_file, size_t ref_len, size_t stride, bool lookup_mode);
std::pair<std::vector<std::string>, std::vector<std::string>> preprocess_fastq(const std::string &fastq_file);
std::pair<std::vector<std::string>, std::vector<std::string>> read_file(const std::string &file_path, size_t ref_len = 150, size_t stride = 1, bool lookup_mode = false);
std::string find_sequence(const std::string &ref_genome, size_t id, size_t ref_len);

// ---- utils
// write_sam (src/utils/utils.cpp:336-403): the reference's writer with its pseudo MAPQ / CIGAR; n_results = final_seqs.size().
void write_sam(size_t n_results, const std::vector<std::string> &query_seqs, const std::vector<std::string> &query_ids,
               const std::vector<size_t> &sequence_ids, const std::string &ref_name, int ref_len, size_t k, const std::string &output_file);
using ConfigValue = std::variant<size_t, float, std::string>;
int save_config(const std::unordered_map<std::string, ConfigValue> &config, const std::string &folder_path, const std::string &config_file = "config.txt");
std::unordered_map<std::string, ConfigValue> load_config(const std::string &config_file);
int save_results(const std::vector<std::vector<size_t>> &neighbors, const std::vector<std::vector<float>> &distances, const std::string &indices_file,
                 const std::string &distances_file, size_t k, const bool use_npy = true);
int save_results(const std::vector<std::vector<long int>> &neighbors, const std::vector<std::vector<float>> &distances, const std::string &indices_file,
                 const std::string &distances_file, size_t k, const bool use_npy = true);

// ---- npy (cnpy-compatible v1.0 files: external/cnpy/cnpy.h:88-131,235-264, external/cnpy/cnpy.cpp:104-155)
namespace npy {
struct Array { std::vector<size_t> shape; size_t word_size = 0; char kind = '?'; bool fortran_order = false; std::vector<char> data; };
Array load(const std::string &path);
void save_raw(const std::string &path, const void *data, char kind, size_t word_size, const std::vector<size_t> &shape);
inline void save(const std::string &p, const float *d, const std::vector<size_t> &s) { save_raw(p, d, 'f', 4, s); }
inline void save(const std::string &p, const size_t *d, const std::vector<size_t> &s) { save_raw(p, d, 'u', 8, s); }
inline void save(const std::string &p, const int *d, const std::vector<size_t> &s) { save_raw(p, d, 'i', 4, s); }
inline void save(const std::string &p, const unsigned int *d, const std::vector<size_t> &s) { save_raw(p, d, 'u', 4, s); }
}  // namespace npy
