// `test_sw_score` — the reference's SW check program (src/test_sw_score.cpp:5-77) on the GPU scorer:
//   test_sw_score <seq1> <seq2>            one pair
//   test_sw_score <sequences.txt> [out]    batch mode: the first 100 consecutive pairs (seq[i], seq[i+1]) of a
//                                          one-sequence-per-line file (tests/test_data_quer.txt), one score per line
//                                          to `out` (default sw_scores_cpp.txt).  The reference hard-codes the paths.
#include <filesystem>
#include <fstream>
#include <iostream>

#include "../../include/drm_b200.h"
#include "drm_host.hpp"

int main(int argc, char *argv[])
{
    try {
        if (argc == 3 && !std::filesystem::exists(argv[1])) {
            std::cout << "SW score: " << calc_sw_score(argv[1], argv[2]) << std::endl;
            return 0;
        }
        if (argc < 2 || argc > 3) {
            std::cerr << "Usage: " << argv[0] << " <seq1> <seq2> | <sequences.txt> [scores_out.txt]" << std::endl;
            return 1;
        }
        const auto seqs = read_file(argv[1]).first;
        if (seqs.size() < 2) throw std::runtime_error("need at least two sequences");
        const size_t n = std::min<size_t>(100, seqs.size() - 1);
        std::string a, b;
        std::vector<int64_t> ao(n + 1, 0), bo(n + 1, 0);
        for (size_t i = 0; i < n; ++i) {
            a += seqs[i]; b += seqs[i + 1];
            ao[i + 1] = (int64_t)a.size(); bo[i + 1] = (int64_t)b.size();
        }
        std::vector<int32_t> scores(n);
        if (drm_sw_pairs(a.data(), ao.data(), b.data(), bo.data(), (int64_t)n, scores.data()) != DRM_OK) throw std::runtime_error(drm_last_error());
        const std::string out = argc == 3 ? argv[2] : "sw_scores_cpp.txt";
        std::ofstream f(out);
        long sum = 0;
        for (size_t i = 0; i < n; ++i) { f << scores[i] << "\n"; sum += scores[i]; }
        std::cout << "Wrote " << n << " scores to " << out << " (sum " << sum << ")" << std::endl;
    } catch (const std::exception &e) {
        std::cerr << "Error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
