// `pipeline` — the reference's CLI (src/main.cpp:10-451) with the search and rerank stages on the GPU.
//
//   pipeline <index_prefix> <query_seqs.fastq|.fasta|.txt|.npy> <ref_seqs.fasta> [EF=128] [K=128] [K_clusters=5]
//            [output_dir=.] [use_dynamic=0] [use_streaming=0]
//
// Same positional arguments, defaults, file layout (<prefix>/<basename>.index, <prefix>/config.txt), error messages and
// outputs (<output_dir>/indices.npy '<u8', distances.npy '<f4' = the RAW search output, src/main.cpp:371-384).
// Differences, all additive:
//   * a sequence query file is embedded on the GPU (Vectorizer over csrc/embed_kernel.cu; model = $DRM_MODEL or the
//     reference's models/finetuned_sgn33-new-a-Apr6.xml relative to the working directory) unless its embeddings are
//     supplied as `<query file>.npy` (or env DRM_QUERY_EMB=<file.npy>); a bare .npy query = the reference's search-only
//     behaviour;
//   * the rerank is the Smith-Waterman variant the reference has implemented but commented out (src/main.cpp:333-341);
//     its result goes to NEW files sw_scores.npy ('<i4', n x K) and sw_indices.npy ('<u8', n x K); DRM_RERANK=none skips it;
//   * DRM_RERANK=l2 runs the reference's ACTIVE post-process instead (post_process_l2_dynamic / _static, src/main.cpp:313-331)
//     over drm_l2_rerank: candidate windows embedded and ranked by L2 distance on the device; result in NEW flat files
//     l2_distances.npy ('<f4') and l2_indices.npy ('<u8'), laid out like the reference's final_dists / final_ids;
//   * an index built with DRM_SHARDS=<W> (config key n_shards; BASELINE.json configs[4]) is searched position-sharded over W
//     GPUs: one drm_shards_search_rerank call (search on every shard, lists exchanged over NVLink - DRM_XCHG=fused|alltoall|
//     allgather - merged and SW-reranked on the GPU that owns the read);
//   * both SW modes rerank from the cleaned genome on the device: use_dynamic=0 applies the static bound rule (number of
//     dense windows), use_dynamic=1 the dynamic one (genome length) - the two rules of find_sequences.
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <iostream>
#include <stdexcept>

#include "drm_host.hpp"

namespace {
using Clock = std::chrono::high_resolution_clock;
long ms_since(Clock::time_point t0) { return (long)std::chrono::duration_cast<std::chrono::milliseconds>(Clock::now() - t0).count(); }
}  // namespace

int main(int argc, char *argv[])
{
    if (argc < 4 || argc > 10) {
        std::cerr << "Usage: " << argv[0]
                  << " <index_prefix> <query_seqs.fastq> <ref_seqs.fasta> [EF] [K] [K_clusters] [output_dir] [use_dynamic] [use_streaming]" << std::endl;
        std::cerr << "  - query input: FASTQ/FASTA/TXT file (with its embeddings in <file>.npy or $DRM_QUERY_EMB) or pre-computed embeddings in .npy format" << std::endl;
        std::cerr << "  - EF: Optional HNSW search parameter (default: " << Config::Search::EF << ")" << std::endl;
        std::cerr << "  - K: Optional number of nearest neighbors to return (default: " << Config::Search::K << ")" << std::endl;
        std::cerr << "  - output_dir: Optional output directory (default: current directory)" << std::endl;
        std::cerr << "  - use_dynamic: Optional flag to load reference sequences dynamically (1) or statically (0). Default: 0" << std::endl;
        std::cerr << "  - use_streaming: with DRM_RERANK=l2 and use_dynamic=1, writes <output_dir>/results.sam instead of the .npy files (1). Default: 0" << std::endl;
        return 1;
    }
    try {
        const auto master_start = Clock::now();
        std::cout << "=== DeepAligner B200 Pipeline ===" << std::endl << std::endl;
        const std::string index_prefix = argv[1];
        const std::string basename = std::filesystem::path(index_prefix).filename().string();
        const std::string index_file = index_prefix + "/" + basename + ".index";
        const std::string config_file = index_prefix + "/" + "config.txt";
        if (!std::filesystem::exists(config_file)) throw std::runtime_error("Config file does not exist: " + config_file);
        auto config = load_config(config_file);
        const size_t ref_len = std::get<size_t>(config["ref_len"]);
        const size_t stride = std::get<size_t>(config["stride"]);
        const std::string query_input_file = argv[2];
        const std::string ref_seqs_file = argv[3];
        const int ef = argc >= 5 ? std::stoi(argv[4]) : Config::Search::EF;
        const int k = argc >= 6 ? std::stoi(argv[5]) : Config::Search::K;
        int k_clusters = Config::Search::K_CLUSTERS;
        if (stride == 1) k_clusters = k;
        else if (argc >= 7) k_clusters = std::stoi(argv[6]);
        const std::string output_dir = argc >= 8 ? argv[7] : ".";
        const bool use_dynamic = argc >= 9 ? std::stoi(argv[8]) != 0 : false;
        const bool use_streaming = argc >= 10 ? std::stoi(argv[9]) != 0 : false;
        const char *rr = getenv("DRM_RERANK");
        const bool rerank = !(rr && !strcmp(rr, "none"));
        // DRM_RERANK=l2: the reference's active post-process (post_process_l2_dynamic / _static, src/main.cpp:313-331) on the
        // device instead of the Smith-Waterman rerank its main keeps commented out (src/main.cpp:333-341)
        const bool rerank_l2 = rr && !strcmp(rr, "l2");

        std::cout << "[MAIN] PIPELINE CONFIG:" << std::endl;
        std::cout << "[MAIN] Input file: " << query_input_file << std::endl;
        std::cout << "[MAIN] Reference file: " << ref_seqs_file << std::endl;

        // ---------------- data loading
        std::cout << "[MAIN] DATA LOADING STEP" << std::endl;
        auto t0 = Clock::now();
        const bool is_npy = std::filesystem::path(query_input_file).extension().string() == ".npy";
        std::vector<std::string> query_sequences, query_ids;
        std::string emb_file = query_input_file;
        // FASTQ queries: reads and genome are parsed on the device and both stages run as one call (DRM_INPUT=host keeps the
        // reference's host parsing and the two separate stages).
        const std::string qext = std::filesystem::path(query_input_file).extension().string();
        const char *input_mode = getenv("DRM_INPUT");
        const bool sharded = config.count("n_shards") && std::holds_alternative<size_t>(config["n_shards"]) && std::get<size_t>(config["n_shards"]) > 1;
        if (sharded && rerank_l2) throw std::runtime_error("DRM_RERANK=l2 is not available on a position-sharded index");
        const bool device_inputs = !sharded && !is_npy && rerank && !rerank_l2 && (qext == ".fastq" || qext == ".fq") && !(input_mode && !strcmp(input_mode, "host"));
        if (!is_npy) {
            if (!device_inputs) {
                std::tie(query_sequences, query_ids) = read_file(query_input_file);
                if (query_sequences.empty()) {
                    std::cerr << "No query_sequences found in input file!" << std::endl;
                    return 1;
                }
            }
            const char *env = getenv("DRM_QUERY_EMB");
            emb_file = env ? env : query_input_file + ".npy";
        }
        npy::Array arr;
        std::vector<float> emb_store;
        const float *emb = nullptr;
        size_t num_queries = 0, dim = 0;
        if (is_npy || std::filesystem::exists(emb_file)) {
            arr = npy::load(emb_file);
            if (arr.shape.size() != 2) throw std::runtime_error("Error: Expected 2D array in .npy file");
            if (arr.kind != 'f' || arr.word_size != 4 || arr.fortran_order) throw std::runtime_error("Error: embeddings must be C-order float32 ('<f4')");
            num_queries = arr.shape[0]; dim = arr.shape[1];
            emb = reinterpret_cast<const float *>(arr.data.data());
            std::cout << "[MAIN] Loaded " << num_queries << " embeddings of dimension " << dim << std::endl;
        } else {
            // the reference's INFERENCE STEP (src/main.cpp:246-268), on the GPU
            const std::string model_path = embedder_model_path();
            if (!std::filesystem::exists(model_path))
                throw std::runtime_error("Embeddings not found: " + emb_file + ", and no embedder model at " + model_path + " (set DRM_MODEL=<model.xml>)");
            std::cout << "[MAIN] INFERENCE STEP" << std::endl;
            const auto ti = Clock::now();
            if (query_sequences.empty()) std::tie(query_sequences, query_ids) = read_file(query_input_file);
            if (query_sequences.empty()) {
                std::cerr << "No query_sequences found in input file!" << std::endl;
                return 1;
            }
            Vectorizer vectorizer(model_path);
            vectorizer.vectorize_flat(query_sequences, emb_store);
            num_queries = query_sequences.size(); dim = Config::Inference::MODEL_OUT_SIZE;
            emb = emb_store.data();
            std::cout << "[MAIN] Inference completed" << std::endl;
            std::cout << "[MAIN] Inference time: " << ms_since(ti) << " ms" << std::endl << std::endl;
        }
        if (!is_npy && !device_inputs && query_sequences.size() != num_queries) throw std::runtime_error("Error: number of embeddings differs from the number of query sequences");
        std::string ref_genome;
        std::vector<std::string> ref_sequences;
        if (rerank_l2 && !is_npy && !use_dynamic) {
            std::cout << "[MAIN] Using STATIC fetching for reference sequences" << std::endl;
            ref_sequences = read_file(ref_seqs_file, ref_len, 1, true).first;
            std::cout << "[MAIN] Loaded " << ref_sequences.size() << " reference sequences from " << ref_seqs_file << std::endl;
        } else if (device_inputs) {
            std::cout << "[MAIN] Reads and reference are parsed on the device (" << (use_dynamic ? "DYNAMIC" : "STATIC") << " bound rule)" << std::endl;
        } else if (!is_npy && rerank) {
            std::cout << "[MAIN] Using " << (use_dynamic ? "DYNAMIC" : "STATIC") << " fetching for reference sequences" << std::endl;
            ref_genome = extract_FASTA_sequence(ref_seqs_file);
        } else {
            std::cout << "[MAIN] Skipping reference loading (not needed for pre-computed embeddings)" << std::endl;
        }
        std::cout << "[MAIN] Total Data loading time: " << ms_since(t0) << " ms" << std::endl << std::endl;

        // ---------------- index
        std::cout << "[MAIN] HNSW INDEX LOADING STEP" << std::endl;
        std::cout << "[MAIN] Index file: " << index_file << std::endl;
        std::cout << "[MAIN] EF: " << ef << std::endl;
        t0 = Clock::now();
        faiss::IndexHNSWPQ *alg_hnsw = nullptr;
        ShardedIndex *alg_shards = nullptr;
        if (sharded) {
            alg_shards = read_index_sharded(index_prefix, config);
            std::cout << "[MAIN] Position-sharded index: " << alg_shards->world << " shards, one per GPU" << std::endl;
        } else {
            if (!std::filesystem::exists(index_file)) throw std::runtime_error("Index file does not exist: " + index_file);
            alg_hnsw = faiss::read_index(index_file.c_str());
        }
        std::cout << "[MAIN] Index loading time: " << ms_since(t0) << " ms" << std::endl << std::endl;

        // ---------------- search
        std::cout << "[MAIN] HNSW SEARCH STEP" << std::endl;
        t0 = Clock::now();
        std::vector<faiss::idx_t> I;
        std::vector<float> D;
        std::vector<int> sw_scores;
        std::vector<size_t> sw_ids;
        std::vector<float> l2_dists;
        std::vector<size_t> l2_ids;
        bool rerank_done = false;
        if (sharded) {
            const bool with_sw = !is_npy && rerank;
            sharded_search_rerank_flat(alg_shards, emb, (int64_t)num_queries, (int)dim, k_clusters, ef, use_dynamic ? 1 : 0, with_sw ? ref_genome : std::string(),
                                       ref_len, query_sequences, stride, (size_t)k, (size_t)k_clusters, I, D, sw_scores, sw_ids);
            rerank_done = with_sw;
            std::cout << "[MAIN] Sharded HNSW search" << (with_sw ? " + SW rerank" : "") << " time (one call): " << ms_since(t0) << " ms" << std::endl << std::endl;
        } else if (device_inputs) {
            stage_device_inputs(alg_hnsw, query_input_file, ref_seqs_file, emb, (int64_t)num_queries,
                                (int)dim, k_clusters, ef, use_dynamic ? 1 : 0, ref_len, stride, (size_t)k, (size_t)k_clusters, I, D, sw_scores, sw_ids);
            std::cout << "[MAIN] HNSW search + SW rerank time (one call, device-side inputs): " << ms_since(t0) << " ms" << std::endl << std::endl;
        } else {
            faiss_search_flat(alg_hnsw, emb, (int64_t)num_queries, (int)dim, k_clusters, ef, I, D);
            std::cout << "[MAIN] HNSW search time: " << ms_since(t0) << " ms" << std::endl << std::endl;
        }
        delete alg_hnsw;
        delete alg_shards;

        // ---------------- post-processing (Smith-Waterman rerank)
        std::cout << "[MAIN] POST-PROCESSING STEP" << std::endl;
        t0 = Clock::now();
        if (device_inputs || rerank_done) {
            std::cout << "[MAIN] Post-processing ran inside the search call" << std::endl;
        } else if (!is_npy && rerank_l2) {
            const std::string model_path = embedder_model_path();
            if (!std::filesystem::exists(model_path)) throw std::runtime_error("DRM_RERANK=l2 needs the embedder model (" + model_path + "; set DRM_MODEL=<model.xml>)");
            Vectorizer vectorizer(model_path);
            post_process_l2_flat(use_dynamic ? 1 : 0, ref_genome, ref_sequences, ref_len, I, D, (int64_t)num_queries, k_clusters, emb, vectorizer, stride,
                                 (size_t)k, (size_t)k_clusters, l2_dists, l2_ids);
            std::cout << "[MAIN] L2 rerank: " << l2_ids.size() << " final candidates" << std::endl;
        } else if (!is_npy && rerank && !sharded) {
            post_process_sw_flat(use_dynamic ? 1 : 0, ref_genome, ref_len, I, (int64_t)num_queries, k_clusters, query_sequences, stride, (size_t)k,
                                 (size_t)k_clusters, sw_scores, sw_ids);
        } else {
            std::cout << "[MAIN] Skipping post-processing (no sequence data available)" << std::endl;
        }
        std::cout << "[MAIN] Post-processing time: " << ms_since(t0) << " ms" << std::endl << std::endl;

        // ---------------- output
        std::cout << "[MAIN] OUTPUT SAVING STEP" << std::endl;
        t0 = Clock::now();
        std::filesystem::create_directories(output_dir);
        // post_process_l2_dynamic_streaming (src/main.cpp:316-320, :408-411): the L2 result goes to results.sam in the reference's
        // write_sam format and the normal output saving is skipped.  (The reference rewrites the file batch by batch; the lines
        // are the same.)
        const bool sam_only = rerank_l2 && use_dynamic && use_streaming && !is_npy;
        if (sam_only) {
            const std::string sam_file = output_dir + "/results.sam";
            std::cout << "[MAIN] Using STREAMING output to SAM file: " << sam_file << std::endl;
            write_sam(l2_ids.size(), query_sequences, query_ids, l2_ids, "ref", (int)ref_len, (size_t)k, sam_file);
            std::cout << "[MAIN] Skip normal output saving since streaming output is used." << std::endl;
        } else {
            std::vector<size_t> u(I.begin(), I.end());  // long -> size_t: -1 becomes 2^64-1 (src/utils/utils.cpp:323-330)
            npy::save(output_dir + "/indices.npy", u.data(), {num_queries, (size_t)k_clusters});
            npy::save(output_dir + "/distances.npy", D.data(), {num_queries, (size_t)k_clusters});
        }
        if (!sw_scores.empty()) {
            npy::save(output_dir + "/sw_scores.npy", sw_scores.data(), {num_queries, (size_t)k});
            npy::save(output_dir + "/sw_indices.npy", sw_ids.data(), {num_queries, (size_t)k});
        }
        if (!l2_ids.empty() && !sam_only) {  // flat, like the reference's final_dists / final_ids (queries without candidates contribute nothing)
            npy::save(output_dir + "/l2_distances.npy", l2_dists.data(), {l2_dists.size()});
            npy::save(output_dir + "/l2_indices.npy", l2_ids.data(), {l2_ids.size()});
        }
        // SAM with real alignments (the reference's write_sam call is commented out, src/main.cpp:395): DRM_SAM=<n> writes the
        // first n reranked windows of every read to <output_dir>/alignments.sam.
        if (const char *sam_n = getenv("DRM_SAM")) {
            const long n_sam = atol(sam_n);
            if (n_sam > 0 && !sw_scores.empty()) {
                if (query_sequences.empty()) std::tie(query_sequences, query_ids) = read_file(query_input_file);
                if (ref_genome.empty()) ref_genome = extract_FASTA_sequence(ref_seqs_file);
                write_sam_aligned(output_dir + "/alignments.sam", "ref", ref_genome, ref_len, query_sequences, query_ids, sw_scores, sw_ids,
                                  (size_t)k, (size_t)n_sam);
                std::cout << "[MAIN] SAM written: " << output_dir << "/alignments.sam (" << n_sam << " alignment(s) per read)" << std::endl;
            }
        }
        std::cout << "[MAIN] Output saving time: " << ms_since(t0) << " ms" << std::endl;
        std::cout << "[MAIN] Finished processing file: " << query_input_file << std::endl;
        std::cout << "[MAIN] Index: " << index_prefix << std::endl;
        std::cout << "[MAIN] Total pipeline time: " << ms_since(master_start) << " ms" << std::endl;
    } catch (const std::exception &e) {
        std::cerr << "Error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
