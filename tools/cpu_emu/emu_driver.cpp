// TEST INFRASTRUCTURE ONLY: ctypes entry points over the emulated kernels (see README.md).
#include "common.h"

#include <string>

using namespace drm;

extern "C" {

const char *emu_last_error() { return last_error_string().c_str(); }

// weights as fp16 bit patterns in the IR's layout; source as EmbedSource describes it (host pointers stand for device ones)
int emu_embed(const uint16_t *emb, int vocab, const uint16_t *W1, const uint16_t *R1, const uint16_t *B1, const uint16_t *W2,
              const uint16_t *R2, const uint16_t *B2, int max_len, int mode, const char *flat, const int64_t *off, const uint8_t *cls,
              int form, int64_t ref_n, int ref_len, uint64_t first_id, uint64_t id_step, const uint64_t *ids, int decorate, int64_t first,
              int64_t n, float *out)
{
    EmbedWeightsHost w;
    w.vocab = vocab;
    auto conv = [](const uint16_t *s, size_t c, std::vector<float> &d) { d.resize(c); for (size_t i = 0; i < c; ++i) d[i] = half_bits_to_float(s[i]); };
    conv(emb, (size_t)vocab * 64, w.emb);
    conv(W1, 2u * 192 * 64, w.W1); conv(R1, 2u * 192 * 64, w.R1); conv(B1, 512, w.B1);
    conv(W2, 2u * 192 * 128, w.W2); conv(R2, 2u * 192 * 64, w.R2); conv(B2, 512, w.B2);
    DeviceEmbedder e;
    int rc = embedder_upload(w, 0, max_len, e);
    if (rc == DRM_OK) {
        EmbedSource src;
        src.mode = mode; src.flat = flat; src.off = off; src.cls = cls; src.form = form; src.ref_n = ref_n; src.ref_len = ref_len;
        src.first_id = first_id; src.id_step = id_step; src.ids = reinterpret_cast<const unsigned long long *>(ids); src.decorate = decorate;
        rc = launch_embed(e, src, first, n, out, nullptr);
    }
    embedder_free(e);
    return rc;
}

int emu_l2_rerank(const float *emb, const uint32_t *map, const int64_t *starts, const float *q, int64_t n, int k, int64_t n_flat,
                  float *out_dist, uint32_t *out_rows, int64_t *counts)
{
    std::vector<float> dist((size_t)n_flat + 1);
    std::vector<int64_t> idx((size_t)n_flat + 1);
    return launch_l2_rerank(emb, map, starts, q, n, k, dist.data(), idx.data(), out_dist, out_rows, counts, nullptr);
}

}  // extern "C"
