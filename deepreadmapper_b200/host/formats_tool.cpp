// `formats_tool` — exercises the host-side format code from the command line, so the CPU tests can compare its
// files byte for byte with what the unmodified reference writes (oracle/_ref).  Not part of the drop-in CLIs.
//   formats_tool save_results <I.bin int64> <D.bin f32> <n> <k_in> <k> <indices.npy> <distances.npy>
//   formats_tool fasta <path> <ref_len> <stride> <lookup 0|1> <windows_out.txt> <labels_out.txt>
//   formats_tool fastq <path> <seqs_out.txt> <ids_out.txt>
//   formats_tool genome <path> <out.txt>
//   formats_tool config <config_in.txt> <folder_out>        (load_config -> save_config)
//   formats_tool npyinfo <file.npy>
//   formats_tool sam <reads.fastq> <ids.bin u64> <n_results> <ref_name> <ref_len> <k> <out.sam>      (write_sam)
//   formats_tool l2flat <mode 0|1> <ref: FASTA (mode 1) | window file for read_file (mode 0)> <ref_len> <I.npy i8> <D.npy f4> <Q.npy f4>
//                <stride> <k> <k_clusters> <dists_out.npy> <ids_out.npy>     (pipeline's DRM_RERANK=l2 stage; model = $DRM_MODEL)
#include <fstream>
#include <iostream>
#include <sstream>

#include "drm_host.hpp"

static std::vector<char> slurp_bin(const std::string &p)
{
    std::ifstream in(p, std::ios::binary);
    if (!in) throw std::runtime_error("cannot open " + p);
    return std::vector<char>((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
}

int main(int argc, char *argv[])
{
    try {
        const std::string cmd = argc > 1 ? argv[1] : "";
        if (cmd == "save_results" && argc == 9) {
            const size_t n = std::stoull(argv[4]), k_in = std::stoull(argv[5]), k = std::stoull(argv[6]);
            const auto ib = slurp_bin(argv[2]), db = slurp_bin(argv[3]);
            const long *I = reinterpret_cast<const long *>(ib.data());
            const float *D = reinterpret_cast<const float *>(db.data());
            std::vector<std::vector<long int>> nb(n);
            std::vector<std::vector<float>> ds(n);
            for (size_t i = 0; i < n; ++i) { nb[i].assign(I + i * k_in, I + (i + 1) * k_in); ds[i].assign(D + i * k_in, D + (i + 1) * k_in); }
            return save_results(nb, ds, argv[7], argv[8], k, true);
        }
        if (cmd == "fasta" && argc == 8) {
            auto [w, l] = preprocess_fasta(argv[2], std::stoull(argv[3]), std::stoull(argv[4]), std::stoi(argv[5]) != 0);
            std::ofstream wo(argv[6]), lo(argv[7]);
            for (const auto &s : w) wo << s << "\n";
            for (size_t v : l) lo << v << "\n";
            return 0;
        }
        if (cmd == "fastq" && argc == 5) {
            auto [s, ids] = preprocess_fastq(argv[2]);
            std::ofstream so(argv[3]), io(argv[4]);
            for (const auto &x : s) so << x << "\n";
            for (const auto &x : ids) io << x << "\n";
            return 0;
        }
        if (cmd == "genome" && argc == 4) {
            std::ofstream o(argv[3]);
            o << extract_FASTA_sequence(argv[2]);
            return 0;
        }
        if (cmd == "config" && argc == 4) return save_config(load_config(argv[2]), argv[3]);
        if (cmd == "npyinfo" && argc == 3) {
            npy::Array a = npy::load(argv[2]);
            std::cout << a.kind << a.word_size << " fortran=" << a.fortran_order << " shape=";
            for (size_t s : a.shape) std::cout << s << ",";
            std::cout << " bytes=" << a.data.size() << std::endl;
            return 0;
        }
        if (cmd == "sam" && argc == 9) {
            auto [seqs, names] = read_file(argv[2]);
            const auto ib = slurp_bin(argv[3]);
            const size_t *ip = reinterpret_cast<const size_t *>(ib.data());
            write_sam(std::stoull(argv[4]), seqs, names, std::vector<size_t>(ip, ip + ib.size() / 8), argv[5], std::stoi(argv[6]), std::stoull(argv[7]), argv[8]);
            return 0;
        }
        if (cmd == "l2flat" && argc == 13) {
            const int mode = std::stoi(argv[2]);
            const size_t ref_len = std::stoull(argv[4]);
            const npy::Array ia = npy::load(argv[5]), da = npy::load(argv[6]), qa = npy::load(argv[7]);
            if (ia.shape.size() != 2 || ia.shape != da.shape || qa.shape.size() != 2 || qa.shape[0] != ia.shape[0] || ia.word_size != 8 || da.word_size != 4)
                throw std::runtime_error("l2flat: I (n x k_in, i8), D (n x k_in, f4), Q (n x 128, f4)");
            const faiss::idx_t *ip = reinterpret_cast<const faiss::idx_t *>(ia.data.data());
            const float *dp = reinterpret_cast<const float *>(da.data.data());
            std::vector<faiss::idx_t> I(ip, ip + ia.shape[0] * ia.shape[1]);
            std::vector<float> D(dp, dp + I.size());
            std::string genome;
            std::vector<std::string> windows;
            if (mode == 1) genome = extract_FASTA_sequence(argv[3]);
            else windows = read_file(argv[3], ref_len, 1, true).first;
            Vectorizer vectorizer(embedder_model_path());
            std::vector<float> dists;
            std::vector<size_t> ids;
            post_process_l2_flat(mode, genome, windows, ref_len, I, D, (int64_t)ia.shape[0], (int64_t)ia.shape[1],
                                 reinterpret_cast<const float *>(qa.data.data()), vectorizer, std::stoull(argv[8]), std::stoull(argv[9]),
                                 std::stoull(argv[10]), dists, ids);
            npy::save(argv[11], dists.data(), {dists.size()});
            npy::save(argv[12], ids.data(), {ids.size()});
            return 0;
        }
        std::cerr << "usage: see the header of formats_tool.cpp" << std::endl;
        return 2;
    } catch (const std::exception &e) {
        std::cerr << "Error: " << e.what() << std::endl;
        return 1;
    }
}
