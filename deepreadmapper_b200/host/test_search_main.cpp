// `test_search` — the reference's search-only driver (src/hnswpq/test_search.cpp:5-134):
//   test_search <index_file> <query.npy> <ef> <k>
// loads the index, runs faiss_search, prints the first 10 x 10 (id, dist) and saves indices.npy as uint32 ('<u4')
// and distances.npy ('<f4') in the current directory, exactly what the reference writes (:112-125).
#include <iostream>

#include "drm_host.hpp"

int main(int argc, char *argv[])
{
    if (argc != 5) {
        std::cerr << "Usage: " << argv[0] << " <index_file> <query_file.npy> <ef> <k>" << std::endl;
        return 1;
    }
    try {
        const int ef = std::stoi(argv[3]), k = std::stoi(argv[4]);
        npy::Array arr = npy::load(argv[2]);
        if (arr.shape.size() != 2) throw std::runtime_error("Error: Expected 2D array in .npy file");
        if (arr.kind != 'f' || arr.word_size != 4) throw std::runtime_error("Error: embeddings must be float32");
        faiss::IndexHNSWPQ *index = faiss::read_index(argv[1]);
        std::vector<faiss::idx_t> I;
        std::vector<float> D;
        const size_t n = arr.shape[0];
        faiss_search_flat(index, reinterpret_cast<const float *>(arr.data.data()), (int64_t)n, (int)arr.shape[1], k, ef, I, D);
        delete index;
        for (size_t i = 0; i < std::min<size_t>(10, n); ++i) {
            std::cout << "Query " << i << ":";
            for (int j = 0; j < std::min(10, k); ++j) std::cout << " (" << I[i * k + j] << ", " << D[i * k + j] << ")";
            std::cout << std::endl;
        }
        std::vector<unsigned int> u(I.begin(), I.end());
        npy::save("indices.npy", u.data(), {n, (size_t)k});
        npy::save("distances.npy", D.data(), {n, (size_t)k});
    } catch (const std::exception &e) {
        std::cerr << "Error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
