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
#include <cstdlib>
#include <algorithm>
#include <filesystem>
#include <iostream>
#include <stdexcept>

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
        std::cout << "[BUILD INDEX] Saving index config to folder: " << index_prefix << std::endl;
        std::unordered_map<std::string, ConfigValue> config = {
            {"index_type", std::string("HNSWPQ")}, {"stride", stride}, {"ref_len", ref_len}, {"n_vects", n}, {"dim", d},
            {"M_hnsw", (size_t)M_hnsw}, {"EFC", (size_t)EFC}, {"M_pq", (size_t)M_pq}, {"nbits", (size_t)nbits}, {"index_file", index_file}};
        save_config(config, index_prefix);
        std::cout << "[BUILD INDEX] Config saved successfully." << std::endl;

        std::cout << "[BUILD INDEX] Building IndexHNSWPQ..." << std::endl;
        const size_t n_train = std::max<size_t>(1, (size_t)((double)n * Config::Build::SAMPLE_RATE));
        std::vector<float> train(n_train * d);
        const double step = (double)n / (double)n_train;
        for (size_t i = 0; i < n_train; ++i) {
            const size_t si = std::min((size_t)((double)i * step), n - 1);
            std::copy(x + si * d, x + (si + 1) * d, train.begin() + (long)(i * d));
        }
        // Builder selection: the GPU builder (csrc/build_gpu.cu) when a device is present and the shape is one it handles,
        // the host OpenMP builder otherwise; DRM_BUILD=cpu|gpu forces one (a forced GPU build fails loudly).
        const char *which = std::getenv("DRM_BUILD");
        const bool gpu_shape = M_hnsw <= 16 && (M_pq == 8 || M_pq == 16) && nbits == 8;
        const bool use_gpu = which ? std::string(which) == "gpu" : (drm_device_count() > 0 && gpu_shape);
        int rc;
        if (use_gpu) {
            double tim[5] = {0, 0, 0, 0, 0};
            rc = drm_build_index_gpu(x, (int64_t)n, (int)d, train.data(), (int64_t)n_train, M_pq, nbits, M_hnsw, EFC, 0, 1234u, index_file.c_str(), tim);
            if (rc == DRM_OK)
                std::cout << "[BUILD INDEX] GPU build: upload " << tim[0] << " s, PQ training " << tim[1] << " s, encoding " << tim[2]
                          << " s, HNSW " << tim[3] << " s, export " << tim[4] << " s" << std::endl;
        } else {
            rc = drm_build_index(x, (int64_t)n, (int)d, train.data(), (int64_t)n_train, M_pq, nbits, M_hnsw, EFC, 0, 1234u, index_file.c_str());
        }
        if (rc != DRM_OK) throw std::runtime_error(drm_last_error());
        std::cout << "[BUILD INDEX] Index saved to: " << index_file << std::endl;
        std::cout << "[BUILD INDEX] Finished index building process." << std::endl;
    } catch (const std::exception &e) {
        std::cerr << "Error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
