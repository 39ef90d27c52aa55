// TEST INFRASTRUCTURE ONLY — never linked into the product (libdrm_cuda.so / pipeline / hnswpq_index).
//
// extern "C" wrappers around the UNMODIFIED reference functions on the SW / post-process half of
// the hot path, compiled from /root/reference in place by oracle/Makefile into
// oracle/_ref/libdrm_ref.so.  tests/ and bench.py's cpu_baseline / --impl reference leg call these
// through ctypes to (a) pin oracle/sw_oracle.c against the real reference and (b) time the
// reference's own CPU code.
//
// Wrapped reference entry points:
//   calc_sw_score            src/utils/metrics.cpp:10-45
//   sw_reranker              src/utils/reranker.cpp:3-51
//   post_process_sw_static   src/utils/post_processor.cpp:454-549
//   post_process_sw_dynamic  src/utils/post_processor.cpp:357-452
//   find_sequence            src/utils/post_processor.cpp:47-66
//   reverse_complement       src/utils/parse_inputs.cpp:43-53
//   preprocess_fasta         src/utils/parse_inputs.cpp:696-721 (format_fasta :223-369)
//   extract_FASTA_sequence   src/utils/parse_inputs.cpp:174-220
//   preprocess_fastq         src/utils/parse_inputs.cpp:1077-1112 (format_fastq :843-950)
//   save_results             src/utils/utils.cpp:264-334
//   save_config/load_config  src/utils/utils.cpp:505-597
//   post_process_l2_static   src/utils/post_processor.cpp:1023-1162 (batch_reranker src/utils/reranker.cpp:98-196)
//   post_process_l2_dynamic  src/utils/post_processor.cpp:552-748
//   calc_l2_dist             src/utils/metrics.cpp:47-61
//   Preprocessor::preprocess src/inference/preprocess.cpp:20-42 (3-mer tokeniser of the embedder)
#include "post_processor.hpp"
#include "preprocess.hpp"
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <unistd.h>
#include <fcntl.h>

// The L2 rerank variants call Vectorizer::vectorize.  OpenVINO is absent, so the embedder behind them is a callback the
// test installs (oracle/embed_oracle.py): everything else those functions do - flattening, find_sequences with
// de-duplication, query boundaries, batch_reranker, calc_l2_dist, partial_sort - is the reference's own code.
extern "C" typedef void (*ref_vectorize_cb)(const char *flat, const int64_t *off, int64_t n, float *out);
static ref_vectorize_cb g_vectorize = nullptr;
std::vector<std::vector<float>> Vectorizer::vectorize(const std::vector<std::string> &input, bool)
{
    if (!g_vectorize) throw std::runtime_error("Vectorizer is not available in the oracle build (no callback installed)");
    std::string flat;
    std::vector<int64_t> off(input.size() + 1, 0);
    for (size_t i = 0; i < input.size(); ++i) { flat += input[i]; off[i + 1] = (int64_t)flat.size(); }
    std::vector<float> out(input.size() * 128);
    g_vectorize(flat.data(), off.data(), (int64_t)input.size(), out.data());
    std::vector<std::vector<float>> res(input.size());
    for (size_t i = 0; i < input.size(); ++i) res[i].assign(out.begin() + (long)(i * 128), out.begin() + (long)((i + 1) * 128));
    return res;
}

namespace {
thread_local std::string g_err;

// The reference prints progress bars and, for stride>1, ~10 lines per query from inside the
// OpenMP loop (post_processor.cpp:107,141-147); silence fd 1 for the duration of a call.
struct Quiet {
    int saved = -1;
    Quiet() {
        fflush(stdout);
        saved = dup(1);
        int nul = open("/dev/null", O_WRONLY);
        if (nul >= 0) { dup2(nul, 1); close(nul); }
    }
    ~Quiet() {
        fflush(stdout);
        std::cout.flush();
        if (saved >= 0) { dup2(saved, 1); close(saved); }
    }
};

std::vector<std::string> unflatten(const char *flat, const int64_t *off, int64_t n)
{
    std::vector<std::string> v((size_t)n);
    for (int64_t i = 0; i < n; ++i) v[(size_t)i].assign(flat + off[i], (size_t)(off[i + 1] - off[i]));
    return v;
}
}  // namespace

extern "C" {

const char *ref_last_error() { return g_err.c_str(); }

int ref_calc_sw_score(const char *a, int64_t la, const char *b, int64_t lb)
{
    return calc_sw_score(std::string(a, (size_t)la), std::string(b, (size_t)lb));
}

// scores[i] = calc_sw_score(a_i, b_i), OpenMP over pairs (used for timing the reference scorer).
void ref_sw_pairs(const char *a_flat, const int64_t *a_off, const char *b_flat, const int64_t *b_off,
                  int64_t n, int32_t *scores, int threads)
{
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
    for (int64_t i = 0; i < n; ++i)
        scores[i] = calc_sw_score(std::string(a_flat + a_off[i], (size_t)(a_off[i + 1] - a_off[i])),
                                  std::string(b_flat + b_off[i], (size_t)(b_off[i + 1] - b_off[i])));
}

int ref_sw_reranker(const char *cand_flat, const int64_t *cand_off, int64_t n_cands, const uint64_t *cand_ids,
                    const char *q, int64_t lq, int64_t k, int32_t *out_scores, uint64_t *out_ids)
{
    try {
        auto cands = unflatten(cand_flat, cand_off, n_cands);
        std::vector<size_t> ids(cand_ids, cand_ids + n_cands);
        auto [seqs, scores, top_ids] = sw_reranker(cands, ids, std::string(q, (size_t)lq), (size_t)k);
        for (size_t i = 0; i < scores.size(); ++i) { out_scores[i] = scores[i]; out_ids[i] = top_ids[i]; }
        return (int)scores.size();
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// neighbors: n x k_in int64 (faiss labels, -1 padded).  mode 0 = static (ref_flat = n_ref windows of
// ref_len bytes), mode 1 = dynamic (ref_flat = genome of n_ref bytes).  Outputs n*k scores / ids.
int ref_post_process_sw(int mode, const int64_t *neighbors, int64_t n, int64_t k_in,
                        const char *ref_flat, int64_t n_ref, int64_t ref_len,
                        const char *q_flat, const int64_t *q_off,
                        int64_t stride, int64_t k, int64_t k_clusters,
                        int32_t *out_scores, uint64_t *out_ids, int64_t *out_count)
{
    try {
        Quiet quiet;
        std::vector<std::vector<long int>> nb((size_t)n);
        std::vector<std::vector<float>> dist((size_t)n);
        for (int64_t i = 0; i < n; ++i) {
            nb[(size_t)i].assign(neighbors + i * k_in, neighbors + (i + 1) * k_in);
            dist[(size_t)i].assign((size_t)k_in, 0.0f);
        }
        auto queries = unflatten(q_flat, q_off, n);
        std::vector<std::string> seqs; std::vector<int> scores; std::vector<size_t> ids;
        if (mode == 0) {
            std::vector<std::string> ref((size_t)n_ref);
            for (int64_t i = 0; i < n_ref; ++i) ref[(size_t)i].assign(ref_flat + i * ref_len, (size_t)ref_len);
            std::tie(seqs, scores, ids) = post_process_sw_static<long int>(nb, dist, ref, queries, (size_t)ref_len,
                                                                          (size_t)stride, (size_t)k, (size_t)k_clusters);
        } else {
            std::string genome(ref_flat, (size_t)n_ref);
            std::tie(seqs, scores, ids) = post_process_sw_dynamic<long int>(nb, dist, genome, queries, (size_t)ref_len,
                                                                           (size_t)stride, (size_t)k, (size_t)k_clusters);
        }
        *out_count = (int64_t)scores.size();
        for (size_t i = 0; i < scores.size(); ++i) { out_scores[i] = scores[i]; out_ids[i] = ids[i]; }
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

int64_t ref_find_sequence(const char *genome, int64_t len, uint64_t id, int64_t ref_len, char *out)
{
    std::string w = find_sequence(std::string(genome, (size_t)len), (size_t)id, (size_t)ref_len);
    memcpy(out, w.data(), w.size());
    return (int64_t)w.size();
}

void ref_reverse_complement(const char *s, int64_t len, char *out)
{
    std::string r = reverse_complement(std::string(s, (size_t)len));
    memcpy(out, r.data(), r.size());
}

// FASTA -> windows.  Call once with out==nullptr to get (n_windows, total_bytes), then again to fill.
int ref_preprocess_fasta(const char *path, int64_t ref_len, int64_t stride, int lookup_mode,
                         char *out_flat, int64_t *out_off, uint64_t *out_labels, int64_t *n_out, int64_t *bytes_out)
{
    try {
        Quiet quiet;
        auto [seqs, labels] = preprocess_fasta(path, (size_t)ref_len, (size_t)stride, lookup_mode != 0);
        int64_t total = 0;
        for (auto &s : seqs) total += (int64_t)s.size();
        *n_out = (int64_t)seqs.size(); *bytes_out = total;
        if (out_flat) {
            int64_t p = 0;
            for (size_t i = 0; i < seqs.size(); ++i) {
                out_off[i] = p; memcpy(out_flat + p, seqs[i].data(), seqs[i].size()); p += (int64_t)seqs[i].size();
                out_labels[i] = labels[i];
            }
            out_off[seqs.size()] = p;
        }
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

int64_t ref_extract_fasta_sequence(const char *path, char *out, int64_t cap)
{
    try {
        Quiet quiet;
        std::string g = extract_FASTA_sequence(path);
        if (out && (int64_t)g.size() <= cap) memcpy(out, g.data(), g.size());
        return (int64_t)g.size();
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

int ref_preprocess_fastq(const char *path, char *seq_flat, int64_t *seq_off, char *id_flat, int64_t *id_off,
                         int64_t *n_out, int64_t *seq_bytes, int64_t *id_bytes)
{
    try {
        Quiet quiet;
        auto [seqs, ids] = preprocess_fastq(path);
        int64_t sb = 0, ib = 0;
        for (auto &s : seqs) sb += (int64_t)s.size();
        for (auto &s : ids) ib += (int64_t)s.size();
        *n_out = (int64_t)seqs.size(); *seq_bytes = sb; *id_bytes = ib;
        if (seq_flat) {
            int64_t p = 0, q = 0;
            for (size_t i = 0; i < seqs.size(); ++i) {
                seq_off[i] = p; memcpy(seq_flat + p, seqs[i].data(), seqs[i].size()); p += (int64_t)seqs[i].size();
            }
            seq_off[seqs.size()] = p;
            for (size_t i = 0; i < ids.size(); ++i) {
                id_off[i] = q; memcpy(id_flat + q, ids[i].data(), ids[i].size()); q += (int64_t)ids[i].size();
            }
            id_off[ids.size()] = q;
        }
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

int ref_save_results(const int64_t *neighbors, const float *distances, int64_t n, int64_t k_in, int64_t k,
                     const char *indices_file, const char *distances_file)
{
    try {
        std::vector<std::vector<long int>> nb((size_t)n);
        std::vector<std::vector<float>> dist((size_t)n);
        for (int64_t i = 0; i < n; ++i) {
            nb[(size_t)i].assign(neighbors + i * k_in, neighbors + (i + 1) * k_in);
            dist[(size_t)i].assign(distances + i * k_in, distances + (i + 1) * k_in);
        }
        return save_results(nb, dist, indices_file, distances_file, (size_t)k, true);
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// write_sam (src/utils/utils.cpp:336-403) on the reads of a FASTQ file parsed by the reference's own read_file.
int ref_write_sam(const char *fastq, const uint64_t *ids, int64_t n_ids, int64_t n_results, const char *ref_name, int ref_len, int64_t k,
                  const char *out_file)
{
    try {
        auto [seqs, names] = read_file(fastq);
        std::vector<std::string> final_seqs((size_t)n_results);
        std::vector<float> final_scores((size_t)n_results, 0.0f);
        std::vector<size_t> sequence_ids(ids, ids + n_ids);
        write_sam(final_seqs, final_scores, seqs, names, sequence_ids, ref_name, ref_len, (size_t)k, out_file);
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// Writes the key set hnswpq_index writes (src/hnswpq/index.cpp:289-302) through the reference's save_config.
int ref_save_config(const char *folder, uint64_t stride, uint64_t ref_len, uint64_t n_vects, uint64_t dim,
                    uint64_t M_hnsw, uint64_t EFC, uint64_t M_pq, uint64_t nbits, const char *index_file)
{
    try {
        std::unordered_map<std::string, ConfigValue> config = {
            {"index_type", std::string("HNSWPQ")}, {"stride", (size_t)stride}, {"ref_len", (size_t)ref_len},
            {"n_vects", (size_t)n_vects}, {"dim", (size_t)dim}, {"M_hnsw", (size_t)M_hnsw}, {"EFC", (size_t)EFC},
            {"M_pq", (size_t)M_pq}, {"nbits", (size_t)nbits}, {"index_file", std::string(index_file)},
        };
        return save_config(config, folder);
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

// Returns the size_t value of `key` through the reference's load_config, or -1 / -2 (missing / not size_t).
int64_t ref_load_config_size(const char *config_file, const char *key)
{
    try {
        auto cfg = load_config(config_file);
        auto it = cfg.find(key);
        if (it == cfg.end()) return -1;
        if (!std::holds_alternative<size_t>(it->second)) return -2;
        return (int64_t)std::get<size_t>(it->second);
    } catch (const std::exception &e) { g_err = e.what(); return -3; }
}

// tokens: n x max_len uint16, rows padded with 0 like Vectorizer::prepareBatch does (src/inference/vectorize.cpp:345-366).
// Only call this with sequences of >= 2 bases from {A,C,G,T,a,c,g,t}: anything else indexes past the reference's
// 96-entry table (src/inference/preprocess.hpp hashToken / preprocess.cpp:28-41).
int ref_tokenize(const char *flat, const int64_t *off, int64_t n, int max_len, uint16_t *tokens)
{
    try {
        Preprocessor pre;
        for (int64_t i = 0; i < n; ++i) {
            const std::string s(flat + off[i], (size_t)(off[i + 1] - off[i]));
            const std::vector<uint16_t> t = pre.preprocess(s, (unsigned)max_len);
            for (int j = 0; j < max_len; ++j) tokens[i * max_len + j] = j < (int)t.size() ? t[(size_t)j] : 0;
        }
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

void ref_set_vectorizer(ref_vectorize_cb cb) { g_vectorize = cb; }

float ref_calc_l2_dist(const float *a, const float *b, int64_t d)
{
    return calc_l2_dist(std::vector<float>(a, a + d), std::vector<float>(b, b + d));
}

// The reference's L2 rerank stage with the installed embedder callback.  mode 0 = static (ref_flat = n_ref strings of
// ref_w bytes each, as read_file returns them), mode 1 = dynamic (ref_flat = genome of n_ref bytes).  q_emb: n x 128.
// Outputs: *out_count values in out_scores / out_ids (n x k when every query has enough candidates).
int ref_post_process_l2(int mode, const int64_t *neighbors, const float *distances, int64_t n, int64_t k_in,
                        const char *ref_flat, int64_t n_ref, int64_t ref_w, int64_t ref_len,
                        const char *q_flat, const int64_t *q_off, const float *q_emb,
                        int64_t stride, int64_t k, int64_t k_clusters,
                        float *out_scores, uint64_t *out_ids, int64_t *out_count)
{
    try {
        Quiet quiet;
        std::vector<std::vector<long int>> nb((size_t)n);
        std::vector<std::vector<float>> dist((size_t)n), qe((size_t)n);
        for (int64_t i = 0; i < n; ++i) {
            nb[(size_t)i].assign(neighbors + i * k_in, neighbors + (i + 1) * k_in);
            dist[(size_t)i].assign(distances + i * k_in, distances + (i + 1) * k_in);
            qe[(size_t)i].assign(q_emb + i * 128, q_emb + (i + 1) * 128);
        }
        auto queries = unflatten(q_flat, q_off, n);
        // the functions only pass the Vectorizer through to Vectorizer::vectorize above, which does not touch `this`
        alignas(Vectorizer) static unsigned char storage[sizeof(Vectorizer)];
        Vectorizer &vec = *reinterpret_cast<Vectorizer *>(storage);
        std::vector<std::string> seqs; std::vector<float> scores; std::vector<size_t> ids;
        if (mode == 0) {
            std::vector<std::string> ref((size_t)n_ref);
            for (int64_t i = 0; i < n_ref; ++i) ref[(size_t)i].assign(ref_flat + i * ref_w, (size_t)ref_w);
            std::tie(seqs, scores, ids) = post_process_l2_static<long int>(nb, dist, ref, queries, (size_t)ref_len, (size_t)stride,
                                                                          (size_t)k, qe, vec, (size_t)k_clusters);
        } else {
            std::string genome(ref_flat, (size_t)n_ref);
            std::tie(seqs, scores, ids) = post_process_l2_dynamic<long int>(nb, dist, genome, queries, (size_t)ref_len, (size_t)stride,
                                                                           (size_t)k, qe, vec, (size_t)k_clusters);
        }
        *out_count = (int64_t)scores.size();
        for (size_t i = 0; i < scores.size(); ++i) { out_scores[i] = scores[i]; out_ids[i] = ids[i]; }
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return -1; }
}

}  // extern "C"
