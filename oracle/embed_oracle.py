"""TEST INFRASTRUCTURE ONLY - CPU restatement (numpy) of the reference's read/window embedder.

Never imported by the product (deepreadmapper_b200/); only tests/, tests/golden/make_embedder_golden.py and
__graft_entry__.smoke() may use it.

What it restates
  * tokenise()   Preprocessor::preprocess          src/inference/preprocess.cpp:20-42, hashToken
                                                   includes/inference/preprocess.hpp:9-52, ids src/utils/tok2index.cpp:3-
  * pad/layout   Vectorizer::prepareBatch          src/inference/vectorize.cpp:345-366 (pad id 0, [MAX_LEN, batch])
  * embed()      the model the reference hands to OpenVINO, models/finetuned_sgn33-new-a-Apr6.xml:
                 Gather(emb.weight 7638x64) -> GRUSequence_154 (bidirectional, hidden 64, linear_before_reset,
                 sigmoid/tanh) -> GRUSequence_198 (input 128) -> [layer-2 forward final state, backward final state]
                 (/Reshape [2,2,B,64] -> /Gather_6 index 1 -> /Gather_7 index 0, /Gather_8 index 1 -> Concat).
                 The sequence lengths the graph feeds both GRUs are the constant 123 (ShapeOf(input)[0] broadcast),
                 so pad tokens are stepped through like any other token and the TopK/ScatterElements pair
                 (pack_padded_sequence's sort / unsort) is the identity.

Parity status: OpenVINO is not in this image, so the reference runtime cannot be run here.  The restatement is pinned
against (a) the reference's own tokeniser compiled in place (oracle/_ref: ref_tokenize), (b) the .onnx weights next to
the IR (same tensors, fp32 vs fp16-compressed) and (c) torch.nn.GRU - the framework the model was exported from -
loaded with those weights (tests/golden/make_embedder_golden.py, tests/test_oracle_embed.py).  Floating point:
fp32 arithmetic on fp16-exact weights; agreement with torch is ~1e-6, the tolerance tests use is 2e-4 absolute on
outputs in (-1, 1).
"""
from __future__ import annotations

import xml.etree.ElementTree as ET

import numpy as np

MAX_LEN = 123  # Config::Inference::MAX_LEN, includes/utils/config.hpp:21
HIDDEN = 64
OUT_DIM = 128
VOCAB = 7638

_VAL = {ord("a"): 0, ord("c"): 1, ord("g"): 2, ord("t"): 3, ord("A"): 0, ord("C"): 1, ord("G"): 2, ord("T"): 3}
_LT, _GT = ord("<"), ord(">")


def token_id(slot: int) -> int:
    """Vocabulary id of hash slot 0..95 (the table in src/utils/tok2index.cpp written as its arithmetic:
    "<xy" -> 7542 + 4x+y; "xy>" -> 7558 + 5(4x+y); "xyz" -> 7559 + 5(4x+y) + z)."""
    if slot < 16:
        return 7542 + slot
    if slot < 32:
        return 7558 + 5 * (slot - 16)
    s = slot - 32
    return 7559 + 5 * (s >> 2) + (s & 3)


def hash_token(t0: int, t1: int, t2: int) -> int:
    """hashToken (includes/inference/preprocess.hpp:33-52) on bytes that already went through tolower; char2Val gives
    7 for anything but a/c/g/t (:9-25), so '<' or '>' in other places, and N, produce hashes the table may not have."""
    v0, v1, v2 = _VAL.get(t0, 7), _VAL.get(t1, 7), _VAL.get(t2, 7)
    if t0 == _LT:
        return (v1 << 2) + v2
    if t2 == _GT:
        return 16 + (v0 << 2) + v1
    return 32 + (v0 << 4) + (v1 << 2) + v2


def tokenise(seq: bytes, max_len: int = MAX_LEN) -> np.ndarray:
    """max_len vocabulary ids, 0-padded; works on the bytes as given - the reference hands the tokeniser its
    '<' + bases + '>' strings (src/main.cpp:249-257, src/utils/parse_inputs.cpp:340,919-921), which makes the first
    token hash('<', '<', b0) = slot 28 + val(b0).  Where the reference is undefined - fewer than 2 bytes (unsigned
    underflow at preprocess.cpp:30) or a hash >= 96 (reads past its 96-entry table) - this restatement and the CUDA
    path emit the pad id 0."""
    out = np.zeros(max_len, dtype=np.uint16)
    n = len(seq)
    ln = min(max_len, n)
    if ln < 2:
        return out
    s = bytes(seq[: ln + 1]).lower()

    def tid(h):
        return token_id(h) if h < 96 else 0

    out[0] = tid(hash_token(_LT, s[0], s[1]))
    for i in range(ln - 2):
        out[i + 1] = tid(hash_token(s[i], s[i + 1], s[i + 2]))
    out[ln - 1] = tid(hash_token(s[ln - 2], s[ln - 1], s[ln] if ln < n else _GT))
    return out


def tokenise_batch(seqs, max_len: int = MAX_LEN) -> np.ndarray:
    return np.stack([tokenise(s, max_len) for s in seqs]) if len(seqs) else np.zeros((0, max_len), np.uint16)


# ---- model files ------------------------------------------------------------------------------------------------
_IR_NAMES = {
    "emb": "emb.weight_compressed",
    "W1": "onnx::GRU_397_compressed", "R1": "onnx::GRU_398_compressed", "B1": "Concat_153_compressed",
    "W2": "onnx::GRU_440_compressed", "R2": "onnx::GRU_441_compressed", "B2": "Concat_197_compressed",
}


def load_ir(xml_path: str, bin_path: str) -> dict:
    """Weights of the OpenVINO IR as fp16 arrays: emb [V,64]; W* [2,192,in]; R* [2,192,64]; B* [2,256].
    Gate order z, r, h; B = [Wb_z+Rb_z, Wb_r+Rb_r, Wb_h, Rb_h] (GRUSequence with linear_before_reset)."""
    blob = open(bin_path, "rb").read()
    consts = {}
    for layer in ET.parse(xml_path).getroot().find("layers"):
        if layer.get("type") == "Const":
            d = layer.find("data")
            consts[layer.get("name")] = (d.get("element_type"), d.get("shape"), int(d.get("offset")), int(d.get("size")))
    w = {}
    for key, name in _IR_NAMES.items():
        et, shape, off, size = consts[name]
        if et != "f16":
            raise ValueError(f"{name}: expected f16, found {et}")
        dims = [int(x) for x in shape.split(",")]
        w[key] = np.frombuffer(blob[off: off + size], dtype=np.float16).reshape(dims).copy()
    return w


def random_weights(seed: int = 0, vocab: int = VOCAB, scale: float = 0.12) -> dict:
    """A model of the same architecture with seeded fp16 weights (for tests that must not depend on the real files)."""
    g = np.random.default_rng(seed)

    def t(*shape):
        return (g.standard_normal(shape) * scale).astype(np.float16)

    return {"emb": t(vocab, 64), "W1": t(2, 192, 64), "R1": t(2, 192, 64), "B1": t(2, 256),
            "W2": t(2, 192, 128), "R2": t(2, 192, 64), "B2": t(2, 256)}


# ---- the network --------------------------------------------------------------------------------------------------
def _sigmoid(x):
    return (1.0 / (1.0 + np.exp(-x, dtype=np.float32))).astype(np.float32)


def _gru_direction(x, W, R, B, reverse: bool):
    """x [n, T, in] fp32 -> (y [n, T, 64], h_final [n, 64]).  One direction of GRUSequence, linear_before_reset:
         z = s(Wz x + Rz h + bz)   r = s(Wr x + Rr h + br)   g = tanh(Wh x + bWh + r * (Rh h + bRh))
         h' = (1 - z) * g + z * h"""
    n, T, _ = x.shape
    H = HIDDEN
    W = W.astype(np.float32); R = R.astype(np.float32); B = B.astype(np.float32)
    xp = x @ W.T  # [n, T, 192]
    h = np.zeros((n, H), np.float32)
    y = np.zeros((n, T, H), np.float32)
    order = range(T - 1, -1, -1) if reverse else range(T)
    for t in order:
        hp = h @ R.T  # [n, 192]
        z = _sigmoid(xp[:, t, 0:H] + hp[:, 0:H] + B[0:H])
        r = _sigmoid(xp[:, t, H:2 * H] + hp[:, H:2 * H] + B[H:2 * H])
        g = np.tanh(xp[:, t, 2 * H:] + B[2 * H:3 * H] + r * (hp[:, 2 * H:] + B[3 * H:4 * H])).astype(np.float32)
        h = ((1.0 - z) * g + z * h).astype(np.float32)
        y[:, t] = h
    return y, h


def embed_tokens(tokens: np.ndarray, w: dict) -> np.ndarray:
    """tokens [n, T] vocabulary ids -> [n, 128] fp32."""
    tokens = np.asarray(tokens).astype(np.int64)
    x = w["emb"].astype(np.float32)[tokens]  # [n, T, 64]
    y1f, _ = _gru_direction(x, w["W1"][0], w["R1"][0], w["B1"][0], False)
    y1b, _ = _gru_direction(x, w["W1"][1], w["R1"][1], w["B1"][1], True)
    x2 = np.concatenate([y1f, y1b], axis=2)
    _, hf = _gru_direction(x2, w["W2"][0], w["R2"][0], w["B2"][0], False)
    _, hb = _gru_direction(x2, w["W2"][1], w["R2"][1], w["B2"][1], True)
    return np.concatenate([hf, hb], axis=1).astype(np.float32)


def embed(seqs, w: dict, max_len: int = MAX_LEN) -> np.ndarray:
    return embed_tokens(tokenise_batch(seqs, max_len), w)
