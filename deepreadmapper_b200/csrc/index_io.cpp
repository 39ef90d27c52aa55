// Reader / writer for the FAISS IndexHNSWPQ on-disk format ("IHNp" wrapping an "IxPq" storage), the
// file the reference's hnswpq_index writes with faiss::write_index (src/hnswpq/index.cpp:188) and its
// pipeline loads with faiss::read_index (src/main.cpp:236).  Layout: SURVEY.md §8c item 1.
#include "common.h"

#include <cerrno>
#include <cstring>

namespace drm {

namespace {

thread_local std::string g_error;

constexpr uint32_t fourcc(char a, char b, char c, char d)
{
    return (uint32_t)(uint8_t)a | ((uint32_t)(uint8_t)b << 8) | ((uint32_t)(uint8_t)c << 16) | ((uint32_t)(uint8_t)d << 24);
}

class In {
  public:
    explicit In(FILE *f) : f_(f) {}
    bool good() const { return good_; }
    template <class T> T scalar()
    {
        T v{};
        if (good_ && fread(&v, sizeof v, 1, f_) != 1) good_ = false;
        return v;
    }
    template <class T> void array(std::vector<T> &v, uint64_t limit)
    {
        uint64_t n = scalar<uint64_t>();
        if (!good_ || n > limit) { good_ = false; return; }
        v.resize((size_t)n);
        if (n && fread(v.data(), sizeof(T), (size_t)n, f_) != (size_t)n) good_ = false;
    }
  private:
    FILE *f_;
    bool good_ = true;
};

class Out {
  public:
    explicit Out(FILE *f) : f_(f) {}
    bool good() const { return good_; }
    template <class T> void scalar(T v) { if (good_ && fwrite(&v, sizeof v, 1, f_) != 1) good_ = false; }
    template <class T> void array(const std::vector<T> &v)
    {
        scalar<uint64_t>((uint64_t)v.size());
        if (good_ && !v.empty() && fwrite(v.data(), sizeof(T), v.size(), f_) != v.size()) good_ = false;
    }
  private:
    FILE *f_;
    bool good_ = true;
};

struct IndexHeader { int32_t d; int64_t ntotal; uint8_t trained; int32_t metric; };

IndexHeader read_header(In &in)
{
    IndexHeader h{};
    h.d = in.scalar<int32_t>();
    h.ntotal = in.scalar<int64_t>();
    (void)in.scalar<int64_t>();  // two legacy dummies (1 << 20)
    (void)in.scalar<int64_t>();
    h.trained = in.scalar<uint8_t>();
    h.metric = in.scalar<int32_t>();
    if (h.metric > 1) (void)in.scalar<float>();  // metric_arg
    return h;
}

void write_header(Out &out, int32_t d, int64_t ntotal)
{
    out.scalar<int32_t>(d);
    out.scalar<int64_t>(ntotal);
    out.scalar<int64_t>((int64_t)1 << 20);
    out.scalar<int64_t>((int64_t)1 << 20);
    out.scalar<uint8_t>(1);
    out.scalar<int32_t>(1);  // METRIC_L2
}

}  // namespace

void set_error(const std::string &msg) { g_error = msg; }
int fail(int code, const std::string &msg) { g_error = msg; return code; }
const std::string &last_error_string() { return g_error; }

int read_index_file(const char *path, HostIndex &ix)
{
    FILE *f = fopen(path, "rb");
    if (!f) return fail(DRM_EIO, std::string("Index file does not exist: ") + path);
    In in(f);
    const uint64_t cap = (uint64_t)1 << 40;
    int rc = DRM_OK;
    do {
        if (in.scalar<uint32_t>() != fourcc('I', 'H', 'N', 'p')) { rc = fail(DRM_EIO, std::string(path) + ": not an IndexHNSWPQ (IHNp) file"); break; }
        IndexHeader h = read_header(in);
        if (!in.good() || h.d <= 0 || h.ntotal < 0) { rc = fail(DRM_EIO, std::string(path) + ": bad index header"); break; }
        if (h.metric != 1) { rc = fail(DRM_EUNSUPPORTED, "only METRIC_L2 IndexHNSWPQ files are supported"); break; }
        ix.d = h.d; ix.ntotal = h.ntotal; ix.metric = h.metric;
        in.array(ix.assign_probas, 4096);
        in.array(ix.cum, 4096);
        in.array(ix.levels, cap);
        in.array(ix.offsets, cap);
        in.array(ix.neighbors, cap);
        ix.entry_point = in.scalar<int32_t>();
        ix.max_level = in.scalar<int32_t>();
        ix.efConstruction = in.scalar<int32_t>();
        ix.efSearch = in.scalar<int32_t>();
        ix.upper_beam = in.scalar<int32_t>();
        if (!in.good()) { rc = fail(DRM_EIO, std::string(path) + ": truncated HNSW block"); break; }
        if (in.scalar<uint32_t>() != fourcc('I', 'x', 'P', 'q')) { rc = fail(DRM_EIO, std::string(path) + ": storage is not IndexPQ (IxPq)"); break; }
        IndexHeader s = read_header(in);
        uint64_t pq_d = in.scalar<uint64_t>();
        ix.M = in.scalar<uint64_t>();
        ix.nbits = in.scalar<uint64_t>();
        if (!in.good() || ix.M == 0 || ix.nbits == 0 || ix.nbits > 16 || pq_d != (uint64_t)ix.d || pq_d % ix.M) {
            rc = fail(DRM_EIO, std::string(path) + ": bad ProductQuantizer header"); break;
        }
        ix.ksub = (uint64_t)1 << ix.nbits;
        ix.dsub = pq_d / ix.M;
        ix.code_size = (ix.nbits * ix.M + 7) / 8;
        in.array(ix.centroids, cap);
        in.array(ix.codes, cap);
        ix.search_type = in.scalar<int32_t>();
        ix.encode_signs = in.scalar<uint8_t>();
        ix.polysemous_ht = in.scalar<int32_t>();
        if (!in.good()) { rc = fail(DRM_EIO, std::string(path) + ": truncated PQ block"); break; }
        // Defensive cross-checks: every length must agree with ntotal / d / M (SURVEY.md §7 hard parts).
        bool ok = s.d == ix.d && s.ntotal == ix.ntotal && (int64_t)ix.levels.size() == ix.ntotal &&
                  (int64_t)ix.offsets.size() == ix.ntotal + 1 && ix.cum.size() >= 2 && ix.cum[0] == 0 &&
                  ix.centroids.size() == ix.M * ix.ksub * ix.dsub &&
                  ix.codes.size() == (uint64_t)ix.ntotal * ix.code_size &&
                  (ix.ntotal == 0 || ix.offsets.back() == ix.neighbors.size());
        if (ok && ix.ntotal > 0) {
            ok = ix.entry_point >= 0 && ix.entry_point < ix.ntotal && ix.max_level >= 0 &&
                 (size_t)ix.max_level + 1 < ix.cum.size() + 0 && ix.levels[(size_t)ix.entry_point] == ix.max_level + 1;
        }
        if (!ok) { rc = fail(DRM_EIO, std::string(path) + ": inconsistent array lengths in index file"); break; }
        // One validation pass over levels, offsets and neighbours: a corrupt (or hostile) file must not turn into
        // out-of-bounds device reads in the search kernel.  Cheap next to the upload.
        const int64_t n = ix.ntotal;
        const size_t n_cum = ix.cum.size();
        bool cum_ok = true;
        for (size_t l = 1; l < n_cum; ++l) cum_ok = cum_ok && ix.cum[l] >= ix.cum[l - 1];
        int bad = cum_ok ? 0 : 1;
#pragma omp parallel for schedule(static) reduction(| : bad)
        for (int64_t v = 0; v < n; ++v) {
            const int32_t lv = ix.levels[(size_t)v];
            if (lv < 1 || (size_t)lv >= n_cum) { bad |= 2; continue; }
            const uint64_t o0 = ix.offsets[(size_t)v], o1 = ix.offsets[(size_t)v + 1];
            if (o1 < o0 || o1 - o0 != (uint64_t)ix.cum[(size_t)lv] || o1 > ix.neighbors.size()) { bad |= 4; continue; }
            for (uint64_t j = o0; j < o1; ++j) {
                const int32_t u = ix.neighbors[j];
                if (u < -1 || u >= n) { bad |= 8; break; }
            }
        }
        if (bad) {
            rc = fail(DRM_EIO, std::string(path) + ": index file fails validation (" + ((bad & 1) ? "cum_nneighbor_per_level not monotone; " : "") +
                                   ((bad & 2) ? "vertex level out of range; " : "") + ((bad & 4) ? "offsets do not match levels; " : "") +
                                   ((bad & 8) ? "neighbour id outside [-1, ntotal); " : "") + ")");
            break;
        }
    } while (false);
    fclose(f);
    return rc;
}

int write_index_file(const char *path, const HostIndex &ix)
{
    FILE *f = fopen(path, "wb");
    if (!f) return fail(DRM_EIO, std::string("cannot create ") + path + ": " + strerror(errno));
    Out out(f);
    out.scalar<uint32_t>(fourcc('I', 'H', 'N', 'p'));
    write_header(out, ix.d, ix.ntotal);
    out.array(ix.assign_probas);
    out.array(ix.cum);
    out.array(ix.levels);
    out.array(ix.offsets);
    out.array(ix.neighbors);
    out.scalar<int32_t>(ix.entry_point);
    out.scalar<int32_t>(ix.max_level);
    out.scalar<int32_t>(ix.efConstruction);
    out.scalar<int32_t>(ix.efSearch);
    out.scalar<int32_t>(ix.upper_beam);
    out.scalar<uint32_t>(fourcc('I', 'x', 'P', 'q'));
    write_header(out, ix.d, ix.ntotal);
    out.scalar<uint64_t>((uint64_t)ix.d);
    out.scalar<uint64_t>(ix.M);
    out.scalar<uint64_t>(ix.nbits);
    out.array(ix.centroids);
    out.array(ix.codes);
    out.scalar<int32_t>(ix.search_type);
    out.scalar<uint8_t>(ix.encode_signs);
    out.scalar<int32_t>(ix.polysemous_ht);
    bool good = out.good();
    if (fclose(f) != 0) good = false;
    if (!good) return fail(DRM_EIO, std::string("short write to ") + path);
    return DRM_OK;
}

}  // namespace drm
