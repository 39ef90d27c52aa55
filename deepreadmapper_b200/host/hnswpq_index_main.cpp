// `hnswpq_index` — the reference's index builder CLI (src/hnswpq/index.cpp:195-316):
//
//   hnswpq_index <ref.(fna|fasta|fa|txt|npy)> <index_prefix> <ref_len> [stride=1] [M_pq=8] [nbits=8] [M_hnsw=16] [EFC=200]
//
// writes <prefix>/<prefix>.index (FAISS "IHNp" layout) and <prefix>/config.txt with the reference's keys
// (index_type, stride, ref_len, n_vects, dim, M_hnsw, EFC, M_pq, nbits, index_file; src/hnswpq/index.cpp:289-304).
// A sequence reference is cut into windows by read_file(ref, ref_len, stride) and embedded on the GPU like the reference
// does with OpenVINO (:266-282; model = $DRM_MODEL or models/finetuned_sgn33-new-a-Apr6.xml), unless its embeddings sit
// beside it as `<ref file>.npy`; a .npy reference is used as-is (:236-265).  Training set = evenly strided
// Config::Build::SAMPLE_RATE sample (create_training_set, :57-84,99).
// The graph is built on the GPU (csrc/build_gpu.cu).  There is no silent CPU fallback: without a device, or for a shape the
// GPU builder does not cover (M_hnsw > 16, M_pq not 8 / 16, nbits != 8), the build FAILS unless DRM_BUILD=cpu asks for the
// host OpenMP builder explicitly.
// DRM_SHARDS=<W> (additive; BASELINE.json configs[4]): the vectors are cut by position into W shards (strand pairs kept
// together), each built as its own graph on GPU g mod #GPUs and written to <prefix>/<prefix>.shard<g>.index with the config
// keys n_shards and shard_<g>_id_lo; `pipeline` then searches them position-sharded.
#include <cstdlib>
#include <algorithm>
#include <filesystem>
#include <iostream>
#include <stdexcept>
#include <thread>

#include "../../include/drm_b200.h"
#include "drm_host.hpp"

int main(int argc, char *argv[])
{
    if (argc < 4 || argc > 9) {
        std::cerr << "Usage: " << argv[0] << " <ref_seq.txt> <index_prefix> <ref_len> [stride] [M_pq] [nbits] [M_hnsw] [EFC]" << std::endl;
        std::cerr << "  stride: step size for sliding window (default: 1)" << std::endl;
        std::cerr << "  M_pq: number of PQ subquantizers (default: 8)" << std::endl;
        std::cerr << "  nbits: bits per subquantizer (default: 8)" << std::endl;
        std::cerr << "  M_hnsw: HNSW connectivity (default: 16)" << std::endl;
        std::cerr << "  EFC: efConstruction parameter (default: 200)" << std::endl;
        return 1;
    }
    try {
        const std::string ref_file = argv[1];
        const std::string index_prefix = argv[2];
        const std::string index_file = index_prefix + "/" + index_prefix + ".index";  // the reference's rule (:212)
        const size_t ref_len = (size_t)std::stoul(argv[3]);
        const size_t stride = argc >= 5 ? (size_t)std::stoul(argv[4]) : 1;
        const int M_pq = argc >= 6 ? std::stoi(argv[5]) : Config::Build::M_PQ;
        const int nbits = argc >= 7 ? std::stoi(argv[6]) : Config::Build::NBITS;
        const int M_hnsw = argc >= 8 ? std::stoi(argv[7]) : Config::Build::M;
        const int EFC = argc >= 9 ? std::stoi(argv[8]) : Config::Build::EFC;

        std::string emb_file = ref_file;
        if (std::filesystem::path(ref_file).extension().string() != ".npy") emb_file = ref_file + ".npy";
        npy::Array arr;
        std::vector<float> emb_store;
        size_t n = 0, d = 0;
        const float *x = nullptr;
        if (std::filesystem::exists(emb_file)) {
            std::cout << "[BUILD INDEX] Loading embeddings directly from .npy file: " << emb_file << std::endl;
            arr = npy::load(emb_file);
            if (arr.shape.size() != 2) {
                std::cerr << "Error: Expected 2D array in .npy file" << std::endl;
                return 1;
            }
            if (arr.kind != 'f' || arr.word_size != 4 || arr.fortran_order) throw std::runtime_error("Error: embeddings must be C-order float32 ('<f4')");
            n = arr.shape[0]; d = arr.shape[1];
            x = reinterpret_cast<const float *>(arr.data.data());
            std::cout << "[BUILD INDEX] Loaded " << n << " embeddings of dimension " << d << std::endl;
        } else {
            const std::string model_path = embedder_model_path();
            if (emb_file == ref_file || !std::filesystem::exists(model_path))
                throw std::runtime_error("Embeddings not found: " + emb_file + ", and no embedder model at " + model_path + " (set DRM_MODEL=<model.xml>)");
            std::cout << "[BUILD INDEX] Reading sequences from file: " << ref_file << std::endl;
            std::vector<std::string> ref_sequences = read_file(ref_file, ref_len, stride).first;
            if (ref_sequences.empty()) throw std::runtime_error("No sequences found in file: " + ref_file);
            std::cout << "[BUILD INDEX] Vectorizing " << ref_sequences.size() << " sequences on the GPU..." << std::endl;
            Vectorizer vectorizer(model_path);
            vectorizer.vectorize_flat(ref_sequences, emb_store);
            n = ref_sequences.size(); d = Config::Inference::MODEL_OUT_SIZE;
            x = emb_store.data();
        }
        if (n == 0) throw std::runtime_error("No sequences found in file: " + ref_file);
        // Builder: the GPU builder; DRM_BUILD=cpu is the explicit opt-in to the host OpenMP builder (no silent fallback)
        const char *which = std::getenv("DRM_BUILD");
        const bool want_cpu = which && std::string(which) == "cpu";
        const bool gpu_shape = M_hnsw <= 16 && (M_pq == 8 || M_pq == 16) && nbits == 8;
        const int n_gpus = drm_device_count();
        if (!want_cpu && n_gpus < 1) throw std::runtime_error("no CUDA device: hnswpq_index builds on the GPU (set DRM_BUILD=cpu for the host OpenMP builder)");
        if (!want_cpu && !gpu_shape)
            throw std::runtime_error("the GPU builder covers M_hnsw <= 16, M_pq in {8, 16}, nbits == 8 (set DRM_BUILD=cpu for the host OpenMP builder)");
        size_t n_shards = 1;
        if (const char *e = std::getenv("DRM_SHARDS")) n_shards = (size_t)std::max(1, std::atoi(e));
        if (n_shards > 16) throw std::runtime_error("DRM_SHARDS: at most 16 position shards");
        // shard g holds ids [lo_g, hi_g): cut at even ids so that a position keeps both strands (id = 2 * position + strand)
        std::vector<size_t> lo(n_shards), hi(n_shards);
        const size_t n_pos = (n + 1) / 2;
        for (size_t g = 0; g < n_shards; ++g) { lo[g] = 2 * (n_pos * g / n_shards); hi[g] = std::min(n, 2 * (n_pos * (g + 1) / n_shards)); }

        std::cout << "[BUILD INDEX] Saving index config to folder: " << index_prefix << std::endl;
        std::unordered_map<std::string, ConfigValue> config = {
            {"index_type", std::string("HNSWPQ")}, {"stride", stride}, {"ref_len", ref_len}, {"n_vects", n}, {"dim", d},
            {"M_hnsw", (size_t)M_hnsw}, {"EFC", (size_t)EFC}, {"M_pq", (size_t)M_pq}, {"nbits", (size_t)nbits}, {"index_file", index_file}};
        if (n_shards > 1) {
            config["n_shards"] = n_shards;
            for (size_t g = 0; g < n_shards; ++g) config["shard_" + std::to_string(g) + "_id_lo"] = lo[g];
        }
        save_config(config, index_prefix);
        std::cout << "[BUILD INDEX] Config saved successfully." << std::endl;

        std::cout << "[BUILD INDEX] Building IndexHNSWPQ" << (n_shards > 1 ? " (" + std::to_string(n_shards) + " position shards)" : std::string()) << "..." << std::endl;
        auto build_one = [&](size_t g, const std::string &out_file, int device, std::string &err) -> int {
            const float *xs = x + lo[g] * d;
            const size_t ns = hi[g] - lo[g];
            const size_t n_train = std::max<size_t>(1, (size_t)((double)ns * Config::Build::SAMPLE_RATE));
            std::vector<float> train(n_train * d);
            const double step = (double)ns / (double)n_train;
            for (size_t i = 0; i < n_train; ++i) {
                const size_t si = std::min((size_t)((double)i * step), ns - 1);
                std::copy(xs + si * d, xs + (si + 1) * d, train.begin() + (long)(i * d));
            }
            int rc;
            if (want_cpu) {
                rc = drm_build_index(xs, (int64_t)ns, (int)d, train.data(), (int64_t)n_train, M_pq, nbits, M_hnsw, EFC, 0, 1234u + (uint32_t)g, out_file.c_str());
            } else {
                double tim[5] = {0, 0, 0, 0, 0};
                rc = drm_build_index_gpu(xs, (int64_t)ns, (int)d, train.data(), (int64_t)n_train, M_pq, nbits, M_hnsw, EFC, device, 1234u + (uint32_t)g, out_file.c_str(), tim);
                if (rc == DRM_OK)
                    std::cout << "[BUILD INDEX] GPU " << device << " build of " << ns << " vectors: upload " << tim[0] << " s, PQ training " << tim[1] << " s, encoding "
                              << tim[2] << " s, HNSW " << tim[3] << " s, export " << tim[4] << " s" << std::endl;
            }
            if (rc != DRM_OK) err = drm_last_error();
            return rc;
        };
        if (n_shards == 1) {
            std::string err;
            if (build_one(0, index_file, 0, err) != DRM_OK) throw std::runtime_error(err);
            std::cout << "[BUILD INDEX] Index saved to: " << index_file << std::endl;
        } else {
            std::vector<int> rcs(n_shards, DRM_OK);
            std::vector<std::string> errs(n_shards);
            const size_t lanes = want_cpu ? 1 : (size_t)std::max(1, n_gpus);
            for (size_t g0 = 0; g0 < n_shards; g0 += lanes) {  // one build per GPU at a time
                std::vector<std::thread> th;
                for (size_t g = g0; g < std::min(n_shards, g0 + lanes); ++g)
                    th.emplace_back([&, g] {
                        const std::string f = index_prefix + "/" + index_prefix + ".shard" + std::to_string(g) + ".index";
                        rcs[g] = build_one(g, f, (int)(g % lanes), errs[g]);
                    });
                for (auto &t : th) t.join();
            }
            for (size_t g = 0; g < n_shards; ++g)
                if (rcs[g] != DRM_OK) throw std::runtime_error(errs[g]);
            std::cout << "[BUILD INDEX] " << n_shards << " shard files saved under: " << index_prefix << std::endl;
        }
        std::cout << "[BUILD INDEX] Finished index building process." << std::endl;
    } catch (const std::exception &e) {
        std::cerr << "Error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
