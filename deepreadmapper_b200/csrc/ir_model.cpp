// Reads the embedder's weights out of an OpenVINO IR pair (.xml graph + .bin blob) - the files the reference passes to
// ov::Core::read_model (src/inference/fast_model.cpp:3-40, Config::Inference::MODEL_PATH includes/utils/config.hpp:18).
// Only the architecture the reference ships is accepted (models/finetuned_sgn33-new-a-Apr6.xml): one embedding Gather
// fed by a [vocab, 64] constant, then two bidirectional GRUSequence layers (hidden 64, linear_before_reset,
// sigmoid/tanh) whose W/R/B inputs are constants.  Tensors are found by graph role (edges into the Gather and the
// GRUSequence ports 3/4/5, through Convert), not by their export-specific names; f16 and f32 constants are accepted.
#include "common.h"

#include <algorithm>
#include <cstring>
#include <fstream>
#include <map>
#include <new>
#include <sstream>
#include <stdexcept>

namespace drm {

float half_bits_to_float(uint16_t h)
{
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu, man = h & 0x3ffu, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {  // subnormal: renormalise
            int e = -1;
            do { man <<= 1; ++e; } while (!(man & 0x400u));
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) bits = sign | 0x7f800000u | (man << 13);
    else bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
}

namespace {

using Attrs = std::map<std::string, std::string>;

// key="value" pairs of one tag body
Attrs parse_attrs(const std::string &tag)
{
    Attrs a;
    size_t i = 0;
    while (i < tag.size()) {
        const size_t eq = tag.find("=\"", i);
        if (eq == std::string::npos) break;
        size_t k = eq;
        while (k > 0 && !isspace((unsigned char)tag[k - 1]) && tag[k - 1] != '<') --k;
        const size_t end = tag.find('"', eq + 2);
        if (end == std::string::npos) break;
        a[tag.substr(k, eq - k)] = tag.substr(eq + 2, end - eq - 2);
        i = end + 1;
    }
    return a;
}

struct Layer { int id = -1; std::string type, name; Attrs data; };

std::vector<int64_t> parse_shape(const std::string &s)
{
    std::vector<int64_t> d;
    std::stringstream ss(s);
    std::string tok;
    while (std::getline(ss, tok, ',')) {
        size_t b = 0;
        while (b < tok.size() && isspace((unsigned char)tok[b])) ++b;
        if (b < tok.size()) d.push_back(std::stoll(tok.substr(b)));
    }
    return d;
}

}  // namespace

static int read_ir_model_impl(const char *xml_path, const char *bin_path, EmbedWeightsHost &w);

// The parser uses std::stoi / stoll / stoull and container .at() on file contents; nothing may be thrown through the C ABI
// (drm_embedder_open_ir, drm_embedder_ir_info), so every exception becomes DRM_EIO here.
int read_ir_model(const char *xml_path, const char *bin_path, EmbedWeightsHost &w)
{
    try {
        return read_ir_model_impl(xml_path, bin_path, w);
    } catch (const std::bad_alloc &) {
        return fail(DRM_EIO, "IR: a tensor size in the .xml does not fit in memory");
    } catch (const std::exception &e) {
        return fail(DRM_EIO, std::string("IR: malformed number or attribute in the .xml (") + e.what() + ")");
    } catch (...) {
        return fail(DRM_EIO, "IR: malformed .xml");
    }
}

static int read_ir_model_impl(const char *xml_path, const char *bin_path, EmbedWeightsHost &w)
{
    std::ifstream xf(xml_path, std::ios::binary);
    if (!xf) return fail(DRM_EIO, std::string("cannot open ") + xml_path);
    std::string xml((std::istreambuf_iterator<char>(xf)), std::istreambuf_iterator<char>());
    std::ifstream bf(bin_path, std::ios::binary);
    if (!bf) return fail(DRM_EIO, std::string("cannot open ") + bin_path);
    std::string blob((std::istreambuf_iterator<char>(bf)), std::istreambuf_iterator<char>());

    std::map<int, Layer> layers;
    for (size_t pos = xml.find("<layer "); pos != std::string::npos; pos = xml.find("<layer ", pos + 1)) {
        const size_t tag_end = xml.find('>', pos);
        if (tag_end == std::string::npos) break;
        const Attrs a = parse_attrs(xml.substr(pos, tag_end - pos));
        Layer L;
        try {
            L.id = std::stoi(a.at("id")); L.type = a.at("type"); L.name = a.count("name") ? a.at("name") : "";
        } catch (...) { return fail(DRM_EIO, "IR: <layer> without id/type"); }
        const bool self_closed = xml[tag_end - 1] == '/';
        if (!self_closed) {
            const size_t body_end = xml.find("</layer>", tag_end);
            const size_t dpos = xml.find("<data", tag_end);
            if (dpos != std::string::npos && body_end != std::string::npos && dpos < body_end) {
                const size_t dend = xml.find('>', dpos);
                if (dend != std::string::npos) L.data = parse_attrs(xml.substr(dpos, dend - dpos));
            }
        }
        layers[L.id] = L;
    }
    std::map<std::pair<int, int>, int> src;  // (to-layer, to-port) -> from-layer
    for (size_t pos = xml.find("<edge "); pos != std::string::npos; pos = xml.find("<edge ", pos + 1)) {
        const size_t tag_end = xml.find('>', pos);
        if (tag_end == std::string::npos) break;
        const Attrs a = parse_attrs(xml.substr(pos, tag_end - pos));
        try {
            src[{std::stoi(a.at("to-layer")), std::stoi(a.at("to-port"))}] = std::stoi(a.at("from-layer"));
        } catch (...) { return fail(DRM_EIO, "IR: malformed <edge>"); }
    }
    if (layers.empty() || src.empty()) return fail(DRM_EIO, "IR: no layers or edges found");

    auto const_behind = [&](int layer, int port) -> const Layer * {
        auto it = src.find({layer, port});
        for (int hops = 0; it != src.end() && hops < 4; ++hops) {
            const Layer &L = layers[it->second];
            if (L.type == "Const") return &L;
            if (L.type != "Convert") return nullptr;
            it = src.find({L.id, 0});
        }
        return nullptr;
    };
    auto load = [&](const Layer *L, const std::vector<int64_t> &want, const char *what, std::vector<float> &out,
                    int64_t *dim0 = nullptr) -> int {
        if (!L) return fail(DRM_EIO, std::string("IR: no constant behind ") + what);
        Attrs::const_iterator et = L->data.find("element_type"), sh = L->data.find("shape"), of = L->data.find("offset"),
                              sz = L->data.find("size");
        if (et == L->data.end() || sh == L->data.end() || of == L->data.end() || sz == L->data.end())
            return fail(DRM_EIO, std::string("IR: incomplete <data> for ") + what);
        std::vector<int64_t> dims = parse_shape(sh->second);
        if (dims.size() != want.size()) return fail(DRM_EUNSUPPORTED, std::string("IR: unexpected rank for ") + what);
        int64_t count = 1;
        for (size_t i = 0; i < dims.size(); ++i) {
            if (want[i] >= 0 && dims[i] != want[i])
                return fail(DRM_EUNSUPPORTED, std::string("IR: unexpected shape for ") + what + " (" + sh->second + ")");
            if (dims[i] <= 0 || dims[i] > ((int64_t)1 << 32) || count > ((int64_t)1 << 40) / dims[i])
                return fail(DRM_EIO, std::string("IR: bad dimension in the shape of ") + what + " (" + sh->second + ")");
            count *= dims[i];
        }
        if (dim0) *dim0 = dims[0];
        const uint64_t off = std::stoull(of->second), size = std::stoull(sz->second);
        const int esz = et->second == "f16" ? 2 : et->second == "f32" ? 4 : 0;
        if (!esz) return fail(DRM_EUNSUPPORTED, std::string("IR: ") + what + " is " + et->second + ", expected f16 or f32");
        if (size != (uint64_t)count * esz || off > blob.size() || size > blob.size() - off)
            return fail(DRM_EIO, std::string("IR: ") + what + " does not fit the .bin file");
        out.resize((size_t)count);
        const char *p = blob.data() + off;
        if (esz == 2)
            for (int64_t i = 0; i < count; ++i) { uint16_t h; std::memcpy(&h, p + 2 * i, 2); out[(size_t)i] = half_bits_to_float(h); }
        else std::memcpy(out.data(), p, (size_t)count * 4);
        return DRM_OK;
    };

    // the embedding: a Gather whose data input is a rank-2 constant and whose index input is the model Parameter
    const Layer *emb = nullptr;
    for (auto &kv : layers) {
        if (kv.second.type != "Gather") continue;
        auto idx = src.find({kv.first, 1});
        if (idx == src.end() || layers[idx->second].type != "Parameter") continue;
        const Layer *c = const_behind(kv.first, 0);
        if (c && parse_shape(c->data.count("shape") ? c->data.at("shape") : "").size() == 2) { emb = c; break; }
    }
    int64_t vocab = 0;
    int rc = load(emb, {-1, 64}, "the token embedding", w.emb, &vocab);
    if (rc != DRM_OK) return rc;
    w.vocab = (int)vocab;

    std::vector<const Layer *> gru;
    for (auto &kv : layers)
        if (kv.second.type == "GRUSequence") gru.push_back(&kv.second);
    if (gru.size() != 2) return fail(DRM_EUNSUPPORTED, "IR: expected exactly two GRUSequence layers, found " + std::to_string(gru.size()));
    for (const Layer *g : gru) {
        auto get = [&](const char *k) { auto it = g->data.find(k); return it == g->data.end() ? std::string() : it->second; };
        if (get("direction") != "bidirectional" || get("hidden_size") != "64" || get("linear_before_reset") != "true")
            return fail(DRM_EUNSUPPORTED, "IR: GRUSequence must be bidirectional, hidden_size 64, linear_before_reset");
        std::string act = get("activations");
        act.erase(std::remove_if(act.begin(), act.end(), [](char c) { return isspace((unsigned char)c); }), act.end());
        if (act != "sigmoid,tanh") return fail(DRM_EUNSUPPORTED, "IR: GRUSequence activations must be sigmoid, tanh");
    }
    if ((rc = load(const_behind(gru[0]->id, 3), {2, 192, 64}, "GRU 1 W", w.W1)) != DRM_OK) return rc;
    if ((rc = load(const_behind(gru[0]->id, 4), {2, 192, 64}, "GRU 1 R", w.R1)) != DRM_OK) return rc;
    if ((rc = load(const_behind(gru[0]->id, 5), {2, 256}, "GRU 1 B", w.B1)) != DRM_OK) return rc;
    if ((rc = load(const_behind(gru[1]->id, 3), {2, 192, 128}, "GRU 2 W", w.W2)) != DRM_OK) return rc;
    if ((rc = load(const_behind(gru[1]->id, 4), {2, 192, 64}, "GRU 2 R", w.R2)) != DRM_OK) return rc;
    if ((rc = load(const_behind(gru[1]->id, 5), {2, 256}, "GRU 2 B", w.B2)) != DRM_OK) return rc;
    return DRM_OK;
}

}  // namespace drm
