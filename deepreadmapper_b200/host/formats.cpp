// On-disk formats of the drop-in stage: .npy (cnpy-compatible), config.txt, FASTA / FASTQ / TXT inputs.
// Behaviour follows the reference's src/utils/utils.cpp and src/utils/parse_inputs.cpp (cited per function);
// the code is written for flat buffers, not translated.
#include "drm_host.hpp"

#include <algorithm>
#include <cctype>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <sstream>
#include <stdexcept>

namespace {

std::string slurp(const std::string &path)
{
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("Could not open file: " + path);
    std::ostringstream ss;
    ss << in.rdbuf();
    return ss.str();
}

}  // namespace

// ------------------------------------------------------------------------------------------ npy
namespace npy {

// Header exactly as cnpy::create_npy_header builds it (external/cnpy/cnpy.h:235-264): the dict is padded with
// spaces so that 10 + len is a multiple of 16 (a full 16 when it already is) and its last byte is '\n'.
void save_raw(const std::string &path, const void *data, char kind, size_t word_size, const std::vector<size_t> &shape)
{
    std::string dict = "{'descr': '<";
    dict += kind;
    dict += std::to_string(word_size);
    dict += "', 'fortran_order': False, 'shape': (";
    dict += std::to_string(shape.at(0));
    for (size_t i = 1; i < shape.size(); ++i) dict += ", " + std::to_string(shape[i]);
    if (shape.size() == 1) dict += ",";
    dict += "), }";
    const size_t remainder = 16 - (10 + dict.size()) % 16;
    dict.append(remainder, ' ');
    dict.back() = '\n';
    std::ofstream out(path, std::ios::binary);
    if (!out) throw std::runtime_error("npy_save: unable to open file " + path);
    const char magic[8] = {(char)0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0};
    out.write(magic, 8);
    const uint16_t hl = (uint16_t)dict.size();
    out.write(reinterpret_cast<const char *>(&hl), 2);
    out.write(dict.data(), (std::streamsize)dict.size());
    size_t count = 1;
    for (size_t s : shape) count *= s;
    out.write(static_cast<const char *>(data), (std::streamsize)(count * word_size));
    if (!out) throw std::runtime_error("npy_save: write failed for " + path);
}

// cnpy::parse_npy_header semantics (external/cnpy/cnpy.cpp:104-155): v1 header, little endian, no dtype check
// beyond recording kind and word size - callers decide what they accept.
Array load(const std::string &path)
{
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("npy_load: Unable to open file " + path);
    char pre[10];
    in.read(pre, 10);
    if (!in || memcmp(pre, "\x93NUMPY", 6) != 0) throw std::runtime_error("npy_load: not an npy file: " + path);
    size_t hl;
    if (pre[6] == 1) {
        hl = (uint8_t)pre[8] | ((size_t)(uint8_t)pre[9] << 8);
    } else {  // v2/v3: 4-byte header length
        char more[2];
        in.read(more, 2);
        hl = (uint8_t)pre[8] | ((size_t)(uint8_t)pre[9] << 8) | ((size_t)(uint8_t)more[0] << 16) | ((size_t)(uint8_t)more[1] << 24);
    }
    std::string h(hl, '\0');
    in.read(h.data(), (std::streamsize)hl);
    if (!in) throw std::runtime_error("npy_load: failed to read header of " + path);
    Array a;
    size_t p = h.find("fortran_order");
    if (p == std::string::npos) throw std::runtime_error("npy_load: failed to find header keyword: 'fortran_order'");
    a.fortran_order = h.compare(p + 16, 4, "True") == 0;
    p = h.find('(');
    size_t e = h.find(')');
    if (p == std::string::npos || e == std::string::npos) throw std::runtime_error("npy_load: failed to find header keyword: '(' or ')'");
    std::string dims = h.substr(p + 1, e - p - 1);
    std::stringstream ds(dims);
    std::string tok;
    while (std::getline(ds, tok, ',')) {
        tok.erase(std::remove_if(tok.begin(), tok.end(), [](unsigned char c) { return std::isspace(c); }), tok.end());
        if (!tok.empty()) a.shape.push_back((size_t)std::stoull(tok));
    }
    p = h.find("descr");
    if (p == std::string::npos) throw std::runtime_error("npy_load: failed to find header keyword: 'descr'");
    p += 9;
    if (h[p] != '<' && h[p] != '|') throw std::runtime_error("npy_load: big-endian arrays are not supported");
    a.kind = h[p + 1];
    a.word_size = (size_t)std::stoul(h.substr(p + 2, h.find('\'', p + 2) - p - 2));
    size_t count = 1;
    for (size_t s : a.shape) count *= s;
    a.data.resize(count * a.word_size);
    in.read(a.data.data(), (std::streamsize)a.data.size());
    if ((size_t)in.gcount() != a.data.size()) throw std::runtime_error("npy_load: failed fread");
    return a;
}

}  // namespace npy

// ------------------------------------------------------------------------------------------ config.txt
// `key: value` lines; values parse as size_t, then float, else string (src/utils/utils.cpp:505-597).
int save_config(const std::unordered_map<std::string, ConfigValue> &config, const std::string &folder_path, const std::string &config_file)
{
    std::filesystem::create_directories(folder_path);
    std::ofstream out(folder_path + "/" + config_file);
    if (!out) throw std::runtime_error("Could not create config file: " + config_file);
    for (const auto &kv : config) {
        out << kv.first << ": ";
        std::visit([&](const auto &v) { out << v; }, kv.second);
        out << "\n";
    }
    return 0;
}

std::unordered_map<std::string, ConfigValue> load_config(const std::string &config_file)
{
    std::ifstream in(config_file);
    if (!in) throw std::runtime_error("Could not open config file: " + config_file);
    std::unordered_map<std::string, ConfigValue> config;
    std::string line;
    auto trim = [](std::string s) {
        const size_t b = s.find_first_not_of(" \t");
        if (b == std::string::npos) return std::string();
        const size_t e = s.find_last_not_of(" \t");
        return s.substr(b, e - b + 1);
    };
    while (std::getline(in, line)) {
        const size_t c = line.find(':');
        if (c == std::string::npos) continue;
        const std::string key = trim(line.substr(0, c)), val = trim(line.substr(c + 1));
        try {
            size_t idx = 0;
            const size_t v = std::stoull(val, &idx);
            if (idx == val.size()) { config[key] = v; continue; }
        } catch (...) {}
        try {
            size_t idx = 0;
            const float v = std::stof(val, &idx);
            if (idx == val.size()) { config[key] = v; continue; }
        } catch (...) {}
        config[key] = val;
    }
    return config;
}

// ------------------------------------------------------------------------------------------ save_results
// indices.npy '<u8' (n, k) - a -1 label becomes 2^64-1 through the long -> size_t conversion - and distances.npy
// '<f4' (n, k) (src/utils/utils.cpp:264-334).
int save_results(const std::vector<std::vector<size_t>> &neighbors, const std::vector<std::vector<float>> &distances, const std::string &indices_file,
                 const std::string &distances_file, size_t k, const bool use_npy)
{
    const size_t n = neighbors.size();
    std::vector<size_t> fi(n * k);
    std::vector<float> fd(n * k);
    for (size_t i = 0; i < n; ++i)
        for (size_t j = 0; j < k; ++j) { fi[i * k + j] = neighbors[i][j]; fd[i * k + j] = distances[i][j]; }
    if (use_npy) {
        npy::save(indices_file, fi.data(), {n, k});
        npy::save(distances_file, fd.data(), {n, k});
        return 0;
    }
    std::ofstream io(indices_file, std::ios::binary);
    if (!io) throw std::runtime_error("Could not create indices file: " + indices_file);
    io.write(reinterpret_cast<const char *>(fi.data()), (std::streamsize)(fi.size() * sizeof(size_t)));
    std::ofstream dout(distances_file, std::ios::binary);
    if (!dout) throw std::runtime_error("Could not create distances file: " + distances_file);
    dout.write(reinterpret_cast<const char *>(fd.data()), (std::streamsize)(fd.size() * sizeof(float)));
    return 0;
}

int save_results(const std::vector<std::vector<long int>> &neighbors, const std::vector<std::vector<float>> &distances, const std::string &indices_file,
                 const std::string &distances_file, size_t k, const bool use_npy)
{
    std::vector<std::vector<size_t>> u(neighbors.size());
    for (size_t i = 0; i < neighbors.size(); ++i) u[i].assign(neighbors[i].begin(), neighbors[i].end());
    return save_results(u, distances, indices_file, distances_file, k, use_npy);
}

// ------------------------------------------------------------------------------------------ sequences
// comp_table of the reference: A<->T, C<->G, N->N, any other byte -> '\0' (src/utils/parse_inputs.cpp:5-14,43-53).
std::string reverse_complement(const std::string &seq)
{
    std::string rc(seq.size(), '\0');
    for (size_t i = 0; i < seq.size(); ++i) {
        char o = 0;
        switch (seq[seq.size() - 1 - i]) {
        case 'A': o = 'T'; break;
        case 'T': o = 'A'; break;
        case 'C': o = 'G'; break;
        case 'G': o = 'C'; break;
        case 'N': o = 'N'; break;
        default: break;
        }
        rc[i] = o;
    }
    return rc;
}

// find_sequence (src/utils/post_processor.cpp:47-66): window id -> (pos = id / 2, strand = id % 2); "" when it overruns.
std::string find_sequence(const std::string &ref_genome, size_t id, size_t ref_len)
{
    const size_t pos = id / 2;
    if (pos + ref_len > ref_genome.size()) return "";
    std::string w = ref_genome.substr(pos, ref_len);
    return (id % 2) ? reverse_complement(w) : w;
}

// The cleaned genome the dynamic mode works on: everything after the first line, upper-cased, only ACGTN kept
// (src/utils/parse_inputs.cpp:174-220 - later '>' header lines are NOT skipped, their ACGTN letters leak in, as in
// the reference; only single-contig references are self-consistent).
std::string extract_FASTA_sequence(const std::string &fasta_file)
{
    const std::string data = slurp(fasta_file);
    size_t p = data.find('\n');
    p = p == std::string::npos ? data.size() : p + 1;
    std::string g;
    g.reserve(data.size() - p);
    for (; p < data.size(); ++p) {
        const char c = (char)std::toupper((unsigned char)data[p]);
        if (c == 'A' || c == 'C' || c == 'G' || c == 'T' || c == 'N') g.push_back(c);
    }
    return g;
}

// format_fasta (src/utils/parse_inputs.cpp:223-369): per contig sliding windows of ref_len every `stride`, forward then
// reverse complement, labels (global_position << 1) | strand with global_position advancing by stride per window;
// lookup_mode = no '<' '>' decoration.
std::pair<std::vector<std::string>, std::vector<size_t>> preprocess_fasta(const std::string &fasta_file, size_t ref_len, size_t stride, bool lookup_mode)
{
    const std::string data = slurp(fasta_file);
    std::vector<std::string> contigs;
    std::string cur;
    bool in_seq = false;
    for (size_t p = 0; p < data.size();) {
        if (data[p] == '>') {
            if (!cur.empty()) { contigs.push_back(std::move(cur)); cur.clear(); }
            while (p < data.size() && data[p] != '\n') ++p;
            if (p < data.size()) ++p;
            in_seq = true;
            continue;
        }
        if (in_seq) {
            const char c = (char)std::toupper((unsigned char)data[p]);
            if (c == 'A' || c == 'C' || c == 'G' || c == 'T' || c == 'N') cur.push_back(c);
        }
        ++p;
    }
    if (!cur.empty()) contigs.push_back(std::move(cur));
    std::vector<std::string> out;
    std::vector<size_t> labels;
    size_t global_position = 0;
    for (const std::string &s : contigs) {
        if (s.size() < ref_len) continue;
        const size_t nw = (s.size() - ref_len) / stride + 1;
        for (size_t i = 0; i < nw; ++i) {
            std::string w = s.substr(i * stride, ref_len), rc = reverse_complement(w);
            if (!lookup_mode) { w = "<" + w + ">"; rc = "<" + rc + ">"; }
            out.push_back(std::move(w));
            out.push_back(std::move(rc));
            labels.push_back((global_position << 1) | 0);
            labels.push_back((global_position << 1) | 1);
            global_position += stride;
        }
    }
    return {std::move(out), std::move(labels)};
}

// format_fastq (src/utils/parse_inputs.cpp:843-950): line 4i+1 is the sequence, returned as '<' + seq + '>' without
// upper-casing; line 4i is the header, id = text after '@' up to the first space, tab or '/'.
std::pair<std::vector<std::string>, std::vector<std::string>> preprocess_fastq(const std::string &fastq_file)
{
    const std::string data = slurp(fastq_file);
    std::vector<std::string> seqs, ids;
    size_t line = 0;
    for (size_t p = 0; p < data.size(); ++line) {
        size_t e = data.find('\n', p);
        if (e == std::string::npos) e = data.size();
        if (line % 4 == 0) {
            size_t b = p;
            if (b < e && data[b] == '@') ++b;
            size_t ie = b;
            while (ie < e && data[ie] != ' ' && data[ie] != '\t' && data[ie] != '/') ++ie;
            ids.emplace_back(data, b, ie - b);
        } else if (line % 4 == 1) {
            std::string s;
            s.reserve(e - p + 2);
            s.push_back('<');
            s.append(data, p, e - p);
            s.push_back('>');
            seqs.push_back(std::move(s));
        }
        p = e + 1;
    }
    return {std::move(seqs), std::move(ids)};
}

// read_file (src/utils/utils.cpp:188-215): dispatch on the extension; .txt = one sequence per line.
std::pair<std::vector<std::string>, std::vector<std::string>> read_file(const std::string &file_path, size_t ref_len, size_t stride, bool lookup_mode)
{
    const std::string ext = std::filesystem::path(file_path).extension().string();
    if (ext == ".fna" || ext == ".fasta" || ext == ".fa") {
        std::cout << "Detected FASTA file format." << std::endl;
        return {preprocess_fasta(file_path, ref_len, stride, lookup_mode).first, {}};
    }
    if (ext == ".fastq" || ext == ".fq") {
        std::cout << "Detected FASTQ file format." << std::endl;
        return preprocess_fastq(file_path);
    }
    if (ext != ".txt") throw std::runtime_error("Unsupported file format: " + ext + ". Only .fna/.fastq/.txt are supported.");
    std::cout << "Detected plain text file format." << std::endl;
    const std::string data = slurp(file_path);
    std::vector<std::string> lines;
    for (size_t p = 0; p < data.size();) {
        size_t e = data.find('\n', p);
        if (e == std::string::npos) e = data.size();
        size_t le = e;
        if (le > p && data[le - 1] == '\r') --le;
        if (le > p) lines.emplace_back(data, p, le - p);
        p = e + 1;
    }
    return {std::move(lines), {}};
}

// ------------------------------------------------------------------------------------------ write_sam
// The reference's SAM writer (src/utils/utils.cpp:336-403; write_sam_streaming :405-503 writes the same lines batch by
// batch): one line per (query, rank) with FLAG = secondary (256) from rank 1 on | reverse strand (16) for odd dense ids,
// POS = id / 2 + 1, and the reference's pseudo fields - MAPQ 60, CIGAR "<read length>M", no mate, QUAL "*".
// `n_results` is final_seqs.size() there: ranks past it are not written.
void write_sam(size_t n_results, const std::vector<std::string> &query_seqs, const std::vector<std::string> &query_ids,
               const std::vector<size_t> &sequence_ids, const std::string &ref_name, int ref_len, size_t k, const std::string &output_file)
{
    std::ofstream out(output_file, std::ios::binary);
    std::string buf = "@HD\tVN:1.0\tSO:unsorted\n@SQ\tSN:" + ref_name + "\tLN:" + std::to_string(ref_len) + "\n";
    for (size_t q = 0; q < query_seqs.size(); ++q) {
        const std::string &raw = query_seqs[q];
        const std::string read = raw.size() > 2 ? raw.substr(1, raw.size() - 2) : raw;  // '<' ... '>' off
        const std::string name = q < query_ids.size() && !query_ids[q].empty() ? query_ids[q] : "S1/" + std::to_string(q + 1) + "/0";
        const std::string tail = "\t60\t" + std::to_string(read.size()) + "M\t*\t0\t0\t" + read + "\t*\n";
        for (size_t r = 0; r < k && q * k + r < n_results; ++r) {
            const size_t id = sequence_ids[q * k + r];
            const int flag = (r ? 256 : 0) | (id & 1 ? 16 : 0);
            buf += name; buf += '\t'; buf += std::to_string(flag); buf += '\t'; buf += ref_name; buf += '\t';
            buf += std::to_string(id / 2 + 1); buf += tail;
        }
        if (buf.size() > (1u << 20)) { out.write(buf.data(), (std::streamsize)buf.size()); buf.clear(); }
    }
    out.write(buf.data(), (std::streamsize)buf.size());
}
