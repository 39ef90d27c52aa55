// The two stages of the hot path behind the reference's own C++ signatures, over the C ABI (include/drm_b200.h).
// No arithmetic of the path happens here: these functions flatten the STL containers the reference passes around,
// call libdrm_cuda.so, and rebuild the containers the reference returns.
#include <filesystem>
#include <variant>
#include "drm_host.hpp"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <sstream>
#include <stdexcept>

#include "../../include/drm_b200.h"

namespace {

[[noreturn]] void throw_last() { throw std::runtime_error(drm_last_error()); }

std::vector<int> device_list()
{
    std::vector<int> devs;
    if (const char *e = getenv("DRM_DEVICES")) {
        std::stringstream ss(e);
        std::string tok;
        while (std::getline(ss, tok, ','))
            if (!tok.empty()) devs.push_back(std::stoi(tok));
    }
    if (devs.empty()) {
        const int n = drm_device_count();
        for (int i = 0; i < n; ++i) devs.push_back(i);
    }
    if (devs.empty()) throw std::runtime_error("no CUDA device: the hot path has no CPU fallback");
    return devs;
}

void flatten(const std::vector<std::string> &v, std::string &flat, std::vector<int64_t> &off)
{
    off.assign(v.size() + 1, 0);
    size_t total = 0;
    for (size_t i = 0; i < v.size(); ++i) { total += v[i].size(); off[i + 1] = (int64_t)total; }
    flat.clear();
    flat.reserve(total + 1);
    for (const auto &s : v) flat += s;
}

// One reference handle is kept alive between calls: the pipeline reranks against the same genome / window table.  The
// cache is keyed on the CONTENT (length + 64-bit hash of every byte), never on an address: a caller may refill a buffer
// in place, or a freed vector may come back at the same address with other windows in it.
uint64_t content_hash(const char *p, size_t n, uint64_t h)
{
    const uint64_t m = 0x9e3779b97f4a7c15ull;
    size_t i = 0;
    for (; i + 8 <= n; i += 8) {
        uint64_t w;
        memcpy(&w, p + i, 8);
        h = (h ^ w) * m;
        h ^= h >> 29;
    }
    uint64_t tail = 0;
    if (i < n) memcpy(&tail, p + i, n - i);
    h = (h ^ tail ^ (uint64_t)n) * m;
    return h ^ (h >> 32);
}

struct RefCache {
    drm_ref *h = nullptr;
    uint64_t hash = 0;
    size_t len = 0, ref_len = 0;
    int form = -1;
    ~RefCache() { if (h) drm_ref_close(h); }
} g_ref;

drm_ref *genome_ref(const std::string &genome, size_t ref_len)
{
    const uint64_t hs = content_hash(genome.data(), genome.size(), 1);
    if (g_ref.h && g_ref.form == 0 && g_ref.hash == hs && g_ref.len == genome.size() && g_ref.ref_len == ref_len) return g_ref.h;
    if (g_ref.h) { drm_ref_close(g_ref.h); g_ref.h = nullptr; }
    const std::vector<int> devs = device_list();
    if (drm_ref_open_genome(genome.data(), (int64_t)genome.size(), (int)ref_len, devs.data(), (int)devs.size(), &g_ref.h) != DRM_OK) throw_last();
    g_ref.form = 0; g_ref.hash = hs; g_ref.len = genome.size(); g_ref.ref_len = ref_len;
    return g_ref.h;
}

// strip = 1: the strings are '<' + window + '>' and the device table holds the bare windows
drm_ref *table_ref(const std::vector<std::string> &ref_seqs, size_t ref_len, size_t strip = 0)
{
    uint64_t hs = 2 + strip;
    for (const auto &s : ref_seqs) {
        if (s.size() != ref_len + 2 * strip || (strip && (s.front() != '<' || s.back() != '>')))
            throw std::runtime_error(strip ? "post_process_l2_static: reference sequences must be '<' + ref_len bases + '>'"
                                           : "post_process_sw_static: reference windows must all be ref_len long");
        hs = content_hash(s.data() + strip, ref_len, hs);
    }
    if (g_ref.h && g_ref.form == 1 && g_ref.hash == hs && g_ref.len == ref_seqs.size() && g_ref.ref_len == ref_len) return g_ref.h;
    if (g_ref.h) { drm_ref_close(g_ref.h); g_ref.h = nullptr; }
    std::string flat;
    flat.reserve(ref_seqs.size() * ref_len);
    for (const auto &s : ref_seqs) flat.append(s, strip, ref_len);
    const std::vector<int> devs = device_list();
    if (drm_ref_open_windows(flat.data(), (int64_t)ref_seqs.size(), (int)ref_len, devs.data(), (int)devs.size(), &g_ref.h) != DRM_OK) throw_last();
    g_ref.form = 1; g_ref.hash = hs; g_ref.len = ref_seqs.size(); g_ref.ref_len = ref_len;
    return g_ref.h;
}

template <typename NeighborType>
void flatten_neighbors(const std::vector<std::vector<NeighborType>> &nb, size_t k_clusters, std::vector<int64_t> &flat, int64_t &k_in)
{
    // post_process_sw_* uses the first min(k_clusters, |row|) ids of every row; rows are padded with -1 (never a valid id:
    // as size_t it is >= every bound, so find_sequences drops it exactly like the reference drops out-of-range ids).
    size_t w = 0;
    for (const auto &r : nb) w = std::max(w, std::min(r.size(), k_clusters));
    k_in = (int64_t)std::max<size_t>(w, 1);
    flat.assign(nb.size() * (size_t)k_in, -1);
    for (size_t i = 0; i < nb.size(); ++i)
        for (size_t j = 0; j < std::min(nb[i].size(), (size_t)k_in); ++j) flat[i * (size_t)k_in + j] = (int64_t)nb[i][j];
}

template <typename NeighborType>
std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> post_process_sw(
    int mode, drm_ref *ref, const std::vector<std::vector<NeighborType>> &neighbors, const std::vector<std::string> &query_seqs, size_t stride,
    size_t k, size_t k_clusters, const std::string *genome, const std::vector<std::string> *table, size_t ref_len)
{
    if (k > k_clusters * 2 * stride)  // src/utils/post_processor.cpp:389-392,486-489
        throw std::runtime_error("Final k too large. Ensure k < k_clusters * 2 * stride to have enough candidates.");
    const size_t n = query_seqs.size();
    if (neighbors.size() < n) throw std::runtime_error("post_process_sw: fewer neighbour rows than queries");
    std::vector<int64_t> nb;
    int64_t k_in = 1;
    std::vector<std::vector<NeighborType>> head(neighbors.begin(), neighbors.begin() + (long)n);
    flatten_neighbors(head, k_clusters, nb, k_in);
    std::string qflat;
    std::vector<int64_t> qoff;
    flatten(query_seqs, qflat, qoff);
    std::vector<int> scores(n * k);
    std::vector<uint64_t> ids(n * k);
    if (n && drm_sw_rerank(ref, mode, nb.data(), (int64_t)n, k_in, qflat.data(), qoff.data(), (int64_t)stride, (int64_t)k, (int64_t)k_clusters,
                           scores.data(), ids.data(), nullptr, nullptr) != DRM_OK)
        throw_last();
    std::vector<std::string> seqs(n * k);
    std::vector<size_t> out_ids(ids.begin(), ids.end());
    for (size_t i = 0; i < n * k; ++i)
        seqs[i] = genome ? find_sequence(*genome, out_ids[i], ref_len) : (out_ids[i] < table->size() ? (*table)[out_ids[i]] : std::string());
    return {std::move(seqs), std::move(scores), std::move(out_ids)};
}

}  // namespace

// ------------------------------------------------------------------------------------------ index + search
namespace faiss {

IndexHNSWPQ::~IndexHNSWPQ() { if (handle) drm_index_close(handle); }

IndexHNSWPQ *read_index(const char *index_file)
{
    const std::vector<int> devs = device_list();
    drm_index *h = nullptr;
    if (drm_index_open(index_file, devs.data(), (int)devs.size(), &h) != DRM_OK) throw_last();
    int64_t info[12];
    drm_index_info(h, info);
    auto *ix = new IndexHNSWPQ;
    ix->handle = h;
    ix->d = (int)info[0];
    ix->ntotal = info[1];
    ix->hnsw.efConstruction = (int)info[7];
    ix->hnsw.efSearch = (int)info[8];
    return ix;
}

}  // namespace faiss

// ------------------------------------------------------------------------------------------ position-sharded index
ShardedIndex::~ShardedIndex() { if (handle) drm_shards_close(handle); }

ShardedIndex *read_index_sharded(const std::string &index_prefix, const std::unordered_map<std::string, std::variant<size_t, float, std::string>> &config)
{
    auto get = [&](const std::string &key) -> size_t {
        auto it = config.find(key);
        if (it == config.end() || !std::holds_alternative<size_t>(it->second)) throw std::runtime_error("config.txt: missing key " + key + " (index not built with DRM_SHARDS)");
        return std::get<size_t>(it->second);
    };
    const size_t W = get("n_shards");
    const std::string base = std::filesystem::path(index_prefix).filename().string();
    std::vector<std::string> files;
    std::vector<int64_t> lo;
    for (size_t g = 0; g < W; ++g) {
        files.push_back(index_prefix + "/" + base + ".shard" + std::to_string(g) + ".index");
        if (!std::filesystem::exists(files.back())) throw std::runtime_error("Index file does not exist: " + files.back());
        lo.push_back((int64_t)get("shard_" + std::to_string(g) + "_id_lo"));
    }
    std::vector<int> devs = device_list();
    if (devs.size() < W) throw std::runtime_error("the index has " + std::to_string(W) + " position shards but only " + std::to_string(devs.size()) + " GPU(s) are visible");
    devs.resize(W);
    std::vector<const char *> cfiles;
    for (auto &f : files) cfiles.push_back(f.c_str());
    drm_shards *h = nullptr;
    if (drm_shards_open(cfiles.data(), lo.data(), devs.data(), (int)W, &h) != DRM_OK) throw_last();
    auto *ix = new ShardedIndex;
    ix->handle = h; ix->world = (int)W; ix->devices = devs;
    return ix;
}

void sharded_search_rerank_flat(ShardedIndex *index, const float *q, int64_t n, int d, int k_search, int ef, int mode, const std::string &genome,
                                size_t ref_len, const std::vector<std::string> &query_seqs, size_t stride, size_t k, size_t k_clusters,
                                std::vector<faiss::idx_t> &I, std::vector<float> &D, std::vector<int> &sw_scores, std::vector<size_t> &sw_ids)
{
    if (!index || !index->handle) throw std::runtime_error("faiss_search: null index");
    if (n <= 0) throw std::runtime_error("Query data is empty");
    int exchange = DRM_XCHG_FUSED;
    if (const char *e = getenv("DRM_XCHG")) {
        const std::string v = e;
        if (v == "allgather") exchange = DRM_XCHG_ALLGATHER;
        else if (v == "alltoall") exchange = DRM_XCHG_ALLTOALL;
        else if (v != "fused") throw std::runtime_error("DRM_XCHG must be allgather, alltoall or fused");
    }
    I.resize((size_t)n * k_search);
    D.resize((size_t)n * k_search);
    const bool rerank = !genome.empty();
    drm_ref *ref = nullptr;
    std::string flat;
    std::vector<int64_t> off;
    if (rerank) {
        if ((int64_t)query_seqs.size() != n) throw std::runtime_error("Error: number of embeddings differs from the number of query sequences");
        if (drm_ref_open_genome(genome.data(), (int64_t)genome.size(), (int)ref_len, index->devices.data(), (int)index->devices.size(), &ref) != DRM_OK) throw_last();
        off.resize((size_t)n + 1, 0);
        size_t total = 0;
        for (auto &sq : query_seqs) total += sq.size();
        flat.reserve(total);
        for (int64_t i = 0; i < n; ++i) { flat += query_seqs[(size_t)i]; off[(size_t)i + 1] = (int64_t)flat.size(); }
        sw_scores.resize((size_t)n * k);
        sw_ids.resize((size_t)n * k);
    }
    static_assert(sizeof(size_t) == sizeof(uint64_t), "size_t is 64-bit");
    const int rc = drm_shards_search_rerank(index->handle, ref, mode, q, n, d, k_search, ef, rerank ? flat.data() : nullptr, rerank ? off.data() : nullptr,
                                            (int64_t)stride, (int64_t)k, (int64_t)k_clusters, exchange, 0, D.data(), I.data(),
                                            rerank ? sw_scores.data() : nullptr, rerank ? reinterpret_cast<uint64_t *>(sw_ids.data()) : nullptr, nullptr);
    if (ref) drm_ref_close(ref);
    if (rc != DRM_OK) throw_last();
}

void faiss_search_flat(faiss::IndexHNSWPQ *index, const float *q, int64_t n, int d, int k, int ef, std::vector<faiss::idx_t> &I, std::vector<float> &D)
{
    if (!index || !index->handle) throw std::runtime_error("faiss_search: null index");
    index->hnsw.efSearch = ef;  // src/hnswpq/search.cpp:13
    if (n <= 0) throw std::runtime_error("Query data is empty");
    I.resize((size_t)n * k);
    D.resize((size_t)n * k);
    if (drm_search(index->handle, q, n, d, k, ef, D.data(), I.data(), nullptr) != DRM_OK) throw_last();
}

std::pair<std::vector<std::vector<faiss::idx_t>>, std::vector<std::vector<float>>> faiss_search(
    faiss::IndexHNSWPQ *index, const std::vector<std::vector<float>> &query_data, int k, int ef)
{
    if (query_data.empty()) throw std::runtime_error("Query data is empty");  // src/hnswpq/search.cpp:16-19
    const size_t n = query_data.size(), d = query_data[0].size();
    std::vector<float> flat(n * d);
    for (size_t i = 0; i < n; ++i) {
        if (query_data[i].size() != d) throw std::runtime_error("faiss_search: ragged query matrix");
        memcpy(&flat[i * d], query_data[i].data(), d * sizeof(float));
    }
    std::vector<faiss::idx_t> I;
    std::vector<float> D;
    faiss_search_flat(index, flat.data(), (int64_t)n, (int)d, k, ef, I, D);
    std::vector<std::vector<faiss::idx_t>> nb(n);
    std::vector<std::vector<float>> ds(n);
    for (size_t i = 0; i < n; ++i) {
        nb[i].assign(I.begin() + (long)(i * k), I.begin() + (long)((i + 1) * k));
        ds[i].assign(D.begin() + (long)(i * k), D.begin() + (long)((i + 1) * k));
    }
    return {std::move(nb), std::move(ds)};
}

// ------------------------------------------------------------------------------------------ Smith-Waterman
int calc_sw_score(const std::string &seq1, const std::string &seq2)
{
    const int64_t a_off[2] = {0, (int64_t)seq1.size()}, b_off[2] = {0, (int64_t)seq2.size()};
    int32_t score = 0;
    if (drm_sw_pairs(seq1.data(), a_off, seq2.data(), b_off, 1, &score) != DRM_OK) throw_last();
    return score;
}

std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> sw_reranker(
    const std::vector<std::string> &cand_seqs, const std::vector<size_t> &cand_ids, const std::string &query_seq, size_t k)
{
    const size_t n = cand_seqs.size();
    if (n < k)  // src/utils/reranker.cpp:26-29
        throw std::runtime_error("Not enough candidates (" + std::to_string(n) + " < " + std::to_string(k) + ")");
    if (n == 0 || k == 0) return {};
    std::string aflat, bflat;
    std::vector<int64_t> aoff, boff(n + 1);
    flatten(cand_seqs, aflat, aoff);
    for (size_t i = 0; i <= n; ++i) boff[i] = (int64_t)(i * query_seq.size());
    for (size_t i = 0; i < n; ++i) bflat += query_seq;
    std::vector<int32_t> scores(n);
    if (drm_sw_pairs(aflat.data(), aoff.data(), bflat.data(), boff.data(), (int64_t)n, scores.data()) != DRM_OK) throw_last();
    // top-k by the same std::partial_sort call as the reference (src/utils/reranker.cpp:35-40): index order, not arithmetic
    std::vector<size_t> idx(n);
    for (size_t i = 0; i < n; ++i) idx[i] = i;
    std::partial_sort(idx.begin(), idx.begin() + (long)k, idx.end(), [&](size_t a, size_t b) { return scores[a] > scores[b]; });
    std::vector<std::string> s(k);
    std::vector<int> sc(k);
    std::vector<size_t> ids(k);
    for (size_t i = 0; i < k; ++i) { s[i] = cand_seqs[idx[i]]; sc[i] = scores[idx[i]]; ids[i] = cand_ids[idx[i]]; }
    return {std::move(s), std::move(sc), std::move(ids)};
}

template <typename NeighborType>
std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> post_process_sw_dynamic(
    const std::vector<std::vector<NeighborType>> &neighbors, const std::vector<std::vector<float>> &, const std::string &ref_genome,
    const std::vector<std::string> &query_seqs, size_t ref_len, size_t stride, size_t k, size_t k_clusters)
{
    return post_process_sw<NeighborType>(1, genome_ref(ref_genome, ref_len), neighbors, query_seqs, stride, k, k_clusters, &ref_genome, nullptr, ref_len);
}

template <typename NeighborType>
std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> post_process_sw_static(
    const std::vector<std::vector<NeighborType>> &neighbors, const std::vector<std::vector<float>> &, const std::vector<std::string> &ref_seqs,
    const std::vector<std::string> &query_seqs, size_t ref_len, size_t stride, size_t k, size_t k_clusters)
{
    return post_process_sw<NeighborType>(0, table_ref(ref_seqs, ref_len), neighbors, query_seqs, stride, k, k_clusters, nullptr, &ref_seqs, ref_len);
}

// explicit instantiations, as src/utils/post_processor.cpp:1169-1175
template std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> post_process_sw_dynamic<size_t>(
    const std::vector<std::vector<size_t>> &, const std::vector<std::vector<float>> &, const std::string &, const std::vector<std::string> &, size_t, size_t, size_t, size_t);
template std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> post_process_sw_dynamic<long int>(
    const std::vector<std::vector<long int>> &, const std::vector<std::vector<float>> &, const std::string &, const std::vector<std::string> &, size_t, size_t, size_t, size_t);
template std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> post_process_sw_static<size_t>(
    const std::vector<std::vector<size_t>> &, const std::vector<std::vector<float>> &, const std::vector<std::string> &, const std::vector<std::string> &, size_t, size_t, size_t, size_t);
template std::tuple<std::vector<std::string>, std::vector<int>, std::vector<size_t>> post_process_sw_static<long int>(
    const std::vector<std::vector<long int>> &, const std::vector<std::vector<float>> &, const std::vector<std::string> &, const std::vector<std::string> &, size_t, size_t, size_t, size_t);

void post_process_sw_flat(int mode, const std::string &genome, size_t ref_len, const std::vector<faiss::idx_t> &I, int64_t n, int64_t k_in,
                          const std::vector<std::string> &query_seqs, size_t stride, size_t k, size_t k_clusters, std::vector<int> &scores,
                          std::vector<size_t> &ids)
{
    if (k > k_clusters * 2 * stride) throw std::runtime_error("Final k too large. Ensure k < k_clusters * 2 * stride to have enough candidates.");
    std::string qflat;
    std::vector<int64_t> qoff;
    flatten(query_seqs, qflat, qoff);
    scores.assign((size_t)n * k, 0);
    std::vector<uint64_t> u((size_t)n * k);
    if (n && drm_sw_rerank(genome_ref(genome, ref_len), mode, I.data(), n, k_in, qflat.data(), qoff.data(), (int64_t)stride, (int64_t)k,
                           (int64_t)k_clusters, scores.data(), u.data(), nullptr, nullptr) != DRM_OK)
        throw_last();
    ids.assign(u.begin(), u.end());
}

// ------------------------------------------------------------------------------------------ device-side inputs
// The search + SW stage with both sequence inputs parsed on the device: the FASTQ file's
// bytes become the resident read set (format_fastq), the FASTA file's bytes the class-coded genome
// (extract_FASTA_sequence), and drm_search_rerank_reads runs both stages with only the embeddings going up.
static std::string slurp_file(const std::string &path)
{
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("Could not open file: " + path);
    std::string data;
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    data.resize(sz > 0 ? (size_t)sz : 0);
    if (sz > 0 && fread(&data[0], 1, (size_t)sz, f) != (size_t)sz) { fclose(f); throw std::runtime_error("Could not read file: " + path); }
    fclose(f);
    return data;
}

int64_t stage_device_inputs(faiss::IndexHNSWPQ *index, const std::string &fastq_file, const std::string &fasta_file, const float *emb,
                            int64_t n_emb, int d, int k_search, int ef, int mode, size_t ref_len, size_t stride, size_t k, size_t k_clusters,
                            std::vector<faiss::idx_t> &I, std::vector<float> &D, std::vector<int> &sw_scores, std::vector<size_t> &sw_ids)
{
    if (!index || !index->handle) throw std::runtime_error("faiss_search: null index");
    if (k > k_clusters * 2 * stride) throw std::runtime_error("Final k too large. Ensure k < k_clusters * 2 * stride to have enough candidates.");
    const std::vector<int> devs = device_list();
    struct Handles {
        drm_reads *reads = nullptr;
        drm_ref *ref = nullptr;
        ~Handles() { if (reads) drm_reads_close(reads); if (ref) drm_ref_close(ref); }
    } h;
    {
        const std::string fq = slurp_file(fastq_file);
        if (drm_reads_open_fastq(fq.data(), (int64_t)fq.size(), devs.data(), (int)devs.size(), &h.reads) != DRM_OK) throw_last();
    }
    int64_t info[4];
    drm_reads_info(h.reads, info);
    if (info[0] == 0) throw std::runtime_error("No query_sequences found in input file!");
    if (info[0] != n_emb) throw std::runtime_error("Error: number of embeddings differs from the number of query sequences");
    {
        const std::string fa = slurp_file(fasta_file);
        int64_t glen = 0;
        if (drm_ref_open_fasta(fa.data(), (int64_t)fa.size(), (int)ref_len, devs.data(), (int)devs.size(), &h.ref, &glen) != DRM_OK) throw_last();
    }
    index->hnsw.efSearch = ef;
    I.resize((size_t)n_emb * k_search);
    D.resize((size_t)n_emb * k_search);
    sw_scores.assign((size_t)n_emb * k, 0);
    std::vector<uint64_t> u((size_t)n_emb * k);
    if (drm_search_rerank_reads(index->handle, h.ref, mode, emb, n_emb, d, k_search, ef, h.reads, (int64_t)stride, (int64_t)k, (int64_t)k_clusters,
                                D.data(), I.data(), sw_scores.data(), u.data(), nullptr, nullptr) != DRM_OK)
        throw_last();
    sw_ids.assign(u.begin(), u.end());
    return info[0];
}

// ------------------------------------------------------------------------------------------ SAM with real alignments
// write_sam (src/utils/utils.cpp:336-398) with its pseudo fields made real: POS, CIGAR and NM
// come from the traceback kernel (drm_sw_align) run on the first `n_out` reranked windows of every read, MAPQ from the gap
// to the best score at another locus.  Same columns and QNAME / FLAG rules as the reference; @SQ LN is the genome length
// (the reference writes the window length there); AS:i / NM:i tags are appended.  A reverse-strand window (odd id) aligns
// the read against the reverse complement, so the line carries the reverse-complemented read, the reversed CIGAR and the
// forward-strand position, as SAM defines for flag 16.
static std::string revcomp(const std::string &s)
{
    std::string r(s.rbegin(), s.rend());
    for (char &c : r) {
        switch (c) {
        case 'A': c = 'T'; break; case 'T': c = 'A'; break; case 'C': c = 'G'; break; case 'G': c = 'C'; break;
        case 'a': c = 't'; break; case 't': c = 'a'; break; case 'c': c = 'g'; break; case 'g': c = 'c'; break;
        default: break;
        }
    }
    return r;
}

void write_sam_aligned(const std::string &sam_file, const std::string &ref_name, const std::string &genome, size_t ref_len,
                       const std::vector<std::string> &query_seqs, const std::vector<std::string> &query_ids, const std::vector<int> &sw_scores,
                       const std::vector<size_t> &sw_ids, size_t k, size_t n_out)
{
    const size_t n = query_seqs.size();
    if (n_out > k) n_out = k;
    if (sw_ids.size() < n * k || sw_scores.size() < n * k) throw std::runtime_error("write_sam_aligned: rerank output is smaller than n x k");
    FILE *f = fopen(sam_file.c_str(), "w");
    if (!f) throw std::runtime_error("Failed to open SAM file: " + sam_file);
    fprintf(f, "@HD\tVN:1.0\tSO:unsorted\n@SQ\tSN:%s\tLN:%zu\n", ref_name.c_str(), genome.size());
    if (n == 0 || n_out == 0) { fclose(f); return; }
    const int max_ops = 64;
    std::vector<uint64_t> ids(n * n_out);
    for (size_t i = 0; i < n; ++i)
        for (size_t j = 0; j < n_out; ++j) ids[i * n_out + j] = (uint64_t)sw_ids[i * k + j];
    std::string qflat;
    std::vector<int64_t> qoff;
    flatten(query_seqs, qflat, qoff);
    const size_t na = n * n_out;
    std::vector<int32_t> score(na), rb(na), re(na), qb(na), qe(na), nops(na), nm(na);
    std::vector<uint32_t> ops(na * max_ops);
    if (drm_sw_align(genome_ref(genome, ref_len), ids.data(), (int64_t)n, (int)n_out, qflat.data(), qoff.data(), 1, max_ops, score.data(), rb.data(),
                     re.data(), qb.data(), qe.data(), nops.data(), nm.data(), ops.data()) != DRM_OK) {
        fclose(f);
        throw_last();
    }
    std::string line;
    for (size_t i = 0; i < n; ++i) {
        std::string clean = query_seqs[i];
        if (clean.size() > 2) clean = clean.substr(1, clean.size() - 2);  // PREFIX / POSTFIX, as write_sam strips them
        const std::string rc_clean = revcomp(clean);
        const std::string qname = (i < query_ids.size() && !query_ids[i].empty()) ? query_ids[i] : "S1/" + std::to_string(i + 1) + "/0";
        for (size_t j = 0; j < n_out; ++j) {
            const size_t a = i * n_out + j;
            const size_t id = sw_ids[i * k + j];
            const size_t pos = id / 2;
            const bool rev = id % 2 == 1;
            int flag = j == 0 ? 0 : 256;
            if (rev) flag |= 16;
            line.clear();
            if (score[a] <= 0 || nops[a] > max_ops) {  // nothing aligned (or an alignment with more than 64 runs): unmapped line
                line = qname + "\t" + std::to_string(flag | 4) + "\t*\t0\t0\t*\t*\t0\t0\t" + clean + "\t*\tAS:i:0\n";
                fputs(line.c_str(), f);
                continue;
            }
            // MAPQ: relative gap to the best reranked score at another locus (no overlap with this window)
            int other = 0;
            for (size_t c = 0; c < k; ++c) {
                const size_t p2 = sw_ids[i * k + c] / 2;
                const size_t dist = p2 > pos ? p2 - pos : pos - p2;
                if (dist >= ref_len && sw_scores[i * k + c] > other) other = sw_scores[i * k + c];
            }
            int mapq = (int)(60.0 * (double)(score[a] - other) / (double)score[a] + 0.5);
            mapq = mapq < 0 ? 0 : (mapq > 60 ? 60 : mapq);
            const int lq = (int)clean.size();
            const int clip_l = rev ? lq - qe[a] : qb[a], clip_r = rev ? qb[a] : lq - qe[a];
            const size_t sam_pos = rev ? pos + ref_len - (size_t)re[a] + 1 : pos + (size_t)rb[a] + 1;
            std::string cigar;
            if (clip_l > 0) cigar += std::to_string(clip_l) + "S";
            for (int x = 0; x < nops[a]; ++x) {
                const uint32_t w = ops[a * max_ops + (rev ? nops[a] - 1 - x : x)];
                cigar += std::to_string(w >> 4) + "MID"[w & 15];
            }
            if (clip_r > 0) cigar += std::to_string(clip_r) + "S";
            line = qname + "\t" + std::to_string(flag) + "\t" + ref_name + "\t" + std::to_string(sam_pos) + "\t" + std::to_string(mapq) + "\t" + cigar +
                   "\t*\t0\t0\t" + (rev ? rc_clean : clean) + "\t*\tAS:i:" + std::to_string(score[a]) + "\tNM:i:" + std::to_string(nm[a]) + "\n";
            fputs(line.c_str(), f);
        }
    }
    fclose(f);
}

// ---------------------------------------------------------------------------------------------------------------------
// Vectorizer (src/inference/vectorize.cpp:4-141) over drm_embedder_open_ir / drm_embed
Vectorizer::Vectorizer(const std::string &model_path, size_t, size_t max_len, size_t model_out_size) : model_out_size_(model_out_size)
{
    if (max_len != Config::Inference::MAX_LEN || model_out_size != Config::Inference::MODEL_OUT_SIZE)
        throw std::runtime_error("Vectorizer: the model takes 123 tokens and gives 128 values");
    std::string bin = model_path;
    const size_t dot = bin.rfind('.');
    if (dot != std::string::npos) bin.erase(dot);
    bin += ".bin";
    if (drm_embedder_open_ir(model_path.c_str(), bin.c_str(), 0, &handle_) != DRM_OK) throw std::runtime_error(drm_last_error());
}

Vectorizer::~Vectorizer() { if (handle_) drm_embedder_close(handle_); }

void Vectorizer::vectorize_flat(const std::vector<std::string> &input, std::vector<float> &out)
{
    std::vector<int64_t> off(input.size() + 1, 0);
    for (size_t i = 0; i < input.size(); ++i) off[i + 1] = off[i] + (int64_t)input[i].size();
    std::string flat;
    flat.reserve((size_t)off.back());
    for (const std::string &s : input) flat += s;
    out.assign(input.size() * model_out_size_, 0.0f);
    if (input.empty()) return;
    if (drm_embed(handle_, flat.data(), off.data(), (int64_t)input.size(), out.data()) != DRM_OK) throw std::runtime_error(drm_last_error());
}

std::vector<std::vector<float>> Vectorizer::vectorize(const std::vector<std::string> &input, bool verbose)
{
    if (verbose) std::cout << "[Inference] Vectorizing " << input.size() << " sequences..." << std::endl;
    std::vector<float> flat;
    vectorize_flat(input, flat);
    std::vector<std::vector<float>> out(input.size());
    for (size_t i = 0; i < input.size(); ++i) out[i].assign(flat.begin() + (long)(i * model_out_size_), flat.begin() + (long)((i + 1) * model_out_size_));
    if (verbose) std::cout << "[Inference] Vectorization completed successfully!" << std::endl;
    return out;
}

std::string embedder_model_path()
{
    const char *env = std::getenv("DRM_MODEL");
    return env ? std::string(env) : std::string(Config::Inference::MODEL_PATH);
}

// ---------------------------------------------------------------------------------------------------------------------
// post_process_l2_dynamic / post_process_l2_static (src/utils/post_processor.cpp:552-748,1023-1162) over drm_l2_rerank
namespace {

template <typename NeighborType>
void l2_inputs(const std::vector<std::vector<NeighborType>> &neighbors, const std::vector<std::vector<float>> &distances,
               const std::vector<std::vector<float>> &query_embeddings, std::vector<int64_t> &nb, std::vector<float> &ds, std::vector<float> &qe,
               int64_t &k_in)
{
    const size_t n = query_embeddings.size();
    if (neighbors.size() != n || distances.size() != n) throw std::runtime_error("post_process_l2: neighbors, distances and query embeddings differ in length");
    k_in = n ? (int64_t)neighbors[0].size() : 0;
    nb.clear(); ds.clear(); qe.clear();
    for (size_t i = 0; i < n; ++i) {
        if ((int64_t)neighbors[i].size() != k_in || (int64_t)distances[i].size() != k_in)
            throw std::runtime_error("post_process_l2: every query must carry the same number of neighbours");
        if (query_embeddings[i].size() != Config::Inference::MODEL_OUT_SIZE) throw std::runtime_error("post_process_l2: query embeddings must have 128 values");
        for (int64_t j = 0; j < k_in; ++j) nb.push_back((int64_t)neighbors[i][(size_t)j]);  // size_t -> long keeps the bit pattern
        ds.insert(ds.end(), distances[i].begin(), distances[i].end());
        qe.insert(qe.end(), query_embeddings[i].begin(), query_embeddings[i].end());
    }
}

}  // namespace

template <typename NeighborType>
std::tuple<std::vector<std::string>, std::vector<float>, std::vector<size_t>> post_process_l2_dynamic(
    const std::vector<std::vector<NeighborType>> &neighbors, const std::vector<std::vector<float>> &distances, const std::string &ref_genome,
    const std::vector<std::string> &, size_t ref_len, size_t stride, size_t k, const std::vector<std::vector<float>> &query_embeddings,
    Vectorizer &vectorizer, size_t k_clusters)
{
    std::vector<int64_t> nb; std::vector<float> ds, qe; int64_t k_in = 0;
    l2_inputs(neighbors, distances, query_embeddings, nb, ds, qe, k_in);
    const int64_t n = (int64_t)query_embeddings.size();
    std::vector<float> scores((size_t)n * std::max(k, k_clusters) + 1);
    std::vector<uint64_t> ids(scores.size());
    int64_t count = 0;
    if (n && drm_l2_rerank(vectorizer.handle(), genome_ref(ref_genome, ref_len), 1, nb.data(), ds.data(), n, k_in, qe.data(), (int64_t)stride, (int64_t)k,
                           (int64_t)k_clusters, scores.data(), ids.data(), &count) != DRM_OK)
        throw_last();
    std::vector<std::string> seqs((size_t)count);
    for (int64_t i = 0; i < count; ++i) seqs[(size_t)i] = find_sequence(ref_genome, (size_t)ids[(size_t)i], ref_len);
    scores.resize((size_t)count);
    return {std::move(seqs), std::move(scores), std::vector<size_t>(ids.begin(), ids.begin() + count)};
}

template <typename NeighborType>
std::tuple<std::vector<std::string>, std::vector<float>, std::vector<size_t>> post_process_l2_static(
    const std::vector<std::vector<NeighborType>> &neighbors, const std::vector<std::vector<float>> &distances,
    const std::vector<std::string> &ref_seqs, const std::vector<std::string> &, size_t ref_len, size_t stride, size_t k,
    const std::vector<std::vector<float>> &query_embeddings, Vectorizer &vectorizer, size_t k_clusters)
{
    // ref_seqs are either the '<' window '>' strings read_file returns (the device table holds the bare windows and the embedder
    // puts the decoration back: device mode 0) or the bare windows of read_file(..., lookup_mode = true), which the reference's
    // main passes (src/main.cpp:190) and embeds as they are: device mode 2
    const bool lookup_windows = !ref_seqs.empty() && ref_seqs[0].size() == ref_len;
    std::vector<int64_t> nb; std::vector<float> ds, qe; int64_t k_in = 0;
    l2_inputs(neighbors, distances, query_embeddings, nb, ds, qe, k_in);
    const int64_t n = (int64_t)query_embeddings.size();
    std::vector<float> scores((size_t)n * std::max(k, k_clusters) + 1);
    std::vector<uint64_t> ids(scores.size());
    int64_t count = 0;
    if (n && drm_l2_rerank(vectorizer.handle(), table_ref(ref_seqs, ref_len, lookup_windows ? 0 : 1), lookup_windows ? 2 : 0, nb.data(), ds.data(), n, k_in, qe.data(), (int64_t)stride, (int64_t)k,
                           (int64_t)k_clusters, scores.data(), ids.data(), &count) != DRM_OK)
        throw_last();
    std::vector<std::string> seqs((size_t)count);
    for (int64_t i = 0; i < count; ++i) seqs[(size_t)i] = ref_seqs[(size_t)ids[(size_t)i]];
    scores.resize((size_t)count);
    return {std::move(seqs), std::move(scores), std::vector<size_t>(ids.begin(), ids.begin() + count)};
}

// Flat form for `pipeline` (DRM_RERANK=l2): the reference's main calls post_process_l2_dynamic / _static right after the
// search (src/main.cpp:313-331).  mode 1 = dynamic over the cleaned genome, mode 0 = static over the window table
// that read_file(ref_seqs_file, ref_len, 1, true) returns (bare windows, embedded undecorated like the reference does).  Output is flat like the reference's final vectors.
void post_process_l2_flat(int mode, const std::string &genome, const std::vector<std::string> &ref_windows, size_t ref_len,
                          const std::vector<faiss::idx_t> &I, const std::vector<float> &D, int64_t n, int64_t k_in, const float *q_emb,
                          Vectorizer &vectorizer, size_t stride, size_t k, size_t k_clusters, std::vector<float> &dists, std::vector<size_t> &ids)
{
    drm_ref *ref = nullptr;
    if (mode == 1) {
        ref = genome_ref(genome, ref_len);
    } else {
        // read_file(..., lookup_mode = true) - what the reference's main passes (src/main.cpp:190) - returns bare windows, which
        // the reference then embeds as they are: device mode 2.  A decorated table ('<' window '>') is device mode 0.
        const bool decorated = !ref_windows.empty() && ref_windows[0].size() == ref_len + 2;
        ref = table_ref(ref_windows, ref_len, decorated ? 1 : 0);
        if (!decorated) mode = 2;
    }
    dists.assign((size_t)n * std::max(k, k_clusters) + 1, 0.0f);
    std::vector<uint64_t> u(dists.size());
    int64_t count = 0;
    if (n && drm_l2_rerank(vectorizer.handle(), ref, mode, I.data(), D.data(), n, k_in, q_emb, (int64_t)stride, (int64_t)k, (int64_t)k_clusters,
                           dists.data(), u.data(), &count) != DRM_OK)
        throw_last();
    dists.resize((size_t)count);
    ids.assign(u.begin(), u.begin() + count);
}

// explicit instantiations, as src/utils/post_processor.cpp:1176-1187
#define DRM_L2_INST(T)                                                                                                                          \
    template std::tuple<std::vector<std::string>, std::vector<float>, std::vector<size_t>> post_process_l2_dynamic<T>(                        \
        const std::vector<std::vector<T>> &, const std::vector<std::vector<float>> &, const std::string &, const std::vector<std::string> &,  \
        size_t, size_t, size_t, const std::vector<std::vector<float>> &, Vectorizer &, size_t);                                                 \
    template std::tuple<std::vector<std::string>, std::vector<float>, std::vector<size_t>> post_process_l2_static<T>(                         \
        const std::vector<std::vector<T>> &, const std::vector<std::vector<float>> &, const std::vector<std::string> &,                        \
        const std::vector<std::string> &, size_t, size_t, size_t, const std::vector<std::vector<float>> &, Vectorizer &, size_t);
DRM_L2_INST(size_t)
DRM_L2_INST(long int)
#undef DRM_L2_INST
