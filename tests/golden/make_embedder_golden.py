"""Generates tests/golden/embedder_golden.npz.  Run in the build container (needs /root/reference):

    python tests/golden/make_embedder_golden.py

What it pins, and what it writes (SURVEY.md §8 f4 - the embedder):
  1. oracle/embed_oracle.tokenise == the reference's own tokeniser compiled in place (oracle/_ref ref_tokenize,
     src/inference/preprocess.cpp:20-42) on random ACGT sequences of many lengths;
  2. the IR weights (models/*.bin, fp16) == the .onnx initialisers (fp32) after fp16 rounding, including the
     GRUSequence bias layout [Wb_z+Rb_z, Wb_r+Rb_r, Wb_h, Rb_h] against ONNX's [Wb_zrh, Rb_zrh];
  3. oracle/embed_oracle.embed_tokens == torch.nn.GRU (2 layers, bidirectional) loaded with those weights.
OpenVINO itself is not in the image, so the reference runtime's own output is NOT part of the pin.
The file holds: the model tensors the test needs (fp16; only the 97 embedding rows the tokeniser can produce, under
`emb_rows` / `emb_ids`), the sequences, their token rows from the reference tokeniser, and the expected embeddings.
"""
import ctypes as C
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import embed_oracle as EO  # noqa: E402
from oracle import bindings as OB  # noqa: E402

REF_MODELS = "/root/reference/models/finetuned_sgn33-new-a-Apr6"


# ---- a minimal protobuf reader: just enough of onnx.ModelProto to list the initialisers -----------------------------
def _varint(b, i):
    v = 0; s = 0
    while True:
        c = b[i]; i += 1
        v |= (c & 0x7F) << s; s += 7
        if c < 0x80:
            return v, i


def _fields(b):
    i = 0
    while i < len(b):
        key, i = _varint(b, i)
        f, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _varint(b, i)
        elif wt == 1:
            v = b[i:i + 8]; i += 8
        elif wt == 2:
            n, i = _varint(b, i); v = b[i:i + n]; i += n
        elif wt == 5:
            v = b[i:i + 4]; i += 4
        else:
            raise ValueError(f"wire type {wt}")
        yield f, wt, v


def onnx_initialisers(path):
    model = open(path, "rb").read()
    graph = next(v for f, wt, v in _fields(model) if f == 7 and wt == 2)
    out = {}
    for f, wt, v in _fields(graph):
        if f != 5 or wt != 2:
            continue
        dims, dtype, name, raw, floats = [], None, None, None, []
        for f2, wt2, v2 in _fields(v):
            if f2 == 1:
                if wt2 == 0:
                    dims.append(v2)
                else:  # packed
                    j = 0
                    while j < len(v2):
                        d, j = _varint(v2, j); dims.append(d)
            elif f2 == 2:
                dtype = v2
            elif f2 == 8:
                name = v2.decode()
            elif f2 == 9:
                raw = v2
            elif f2 == 4:
                floats.append(v2 if wt2 == 2 else struct.pack("<f", struct.unpack("<f", v2)[0]))
        if dtype == 1:
            buf = raw if raw is not None else b"".join(floats)
            out[name] = np.frombuffer(buf, dtype="<f4").reshape(dims)
    return out


def main():
    w = EO.load_ir(REF_MODELS + ".xml", REF_MODELS + ".bin")
    for k, v in w.items():
        print(k, v.shape, v.dtype)

    # 2. IR vs ONNX
    ini = onnx_initialisers(REF_MODELS + ".onnx")
    by_shape = {}
    for name, a in ini.items():
        by_shape.setdefault(a.shape, []).append(name)
    print({k: v for k, v in by_shape.items()})
    assert np.array_equal(ini["emb.weight"].astype(np.float16), w["emb"])
    onnx_W1 = next(a for n, a in ini.items() if a.shape == (2, 192, 64) and np.array_equal(a.astype(np.float16), w["W1"]))
    onnx_R1 = next(a for n, a in ini.items() if a.shape == (2, 192, 64) and np.array_equal(a.astype(np.float16), w["R1"]))
    onnx_W2 = next(a for n, a in ini.items() if a.shape == (2, 192, 128) and np.array_equal(a.astype(np.float16), w["W2"]))
    onnx_R2 = next(a for n, a in ini.items() if a.shape == (2, 192, 64) and np.array_equal(a.astype(np.float16), w["R2"]))
    biases = [a for n, a in ini.items() if a.shape == (2, 384)]
    assert len(biases) == 2, [a.shape for a in ini.values()]

    def fold(b):  # ONNX [Wb_z, Wb_r, Wb_h, Rb_z, Rb_r, Rb_h] -> GRUSequence [Wb_z+Rb_z, Wb_r+Rb_r, Wb_h, Rb_h]
        H = 64
        return np.concatenate([b[:, 0:H] + b[:, 3 * H:4 * H], b[:, H:2 * H] + b[:, 4 * H:5 * H], b[:, 2 * H:3 * H], b[:, 5 * H:6 * H]], axis=1)

    folded = [fold(b).astype(np.float16) for b in biases]
    i1 = next(i for i, f in enumerate(folded) if np.array_equal(f, w["B1"]))
    i2 = next(i for i, f in enumerate(folded) if np.array_equal(f, w["B2"]))
    assert {i1, i2} == {0, 1}
    print("IR weights == ONNX initialisers (fp16-rounded), bias layout confirmed")

    # 1. tokeniser vs the reference's
    g = np.random.default_rng(20240611)
    seqs = []
    for ln in list(range(2, 12)) + [60, 100, 121, 122, 123, 124, 125, 150, 150, 150, 150, 151, 200, 300]:
        for _ in range(8):
            s = bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), ln))
            if g.random() < 0.25:
                s = s.lower()
            if g.random() < 0.5:  # the form the reference feeds: PREFIX + bases + POSTFIX
                s = b"<" + s + b">"
            seqs.append(s)
    flat = b"".join(seqs)
    off = np.zeros(len(seqs) + 1, np.int64); off[1:] = np.cumsum([len(s) for s in seqs])
    ref_tok = np.zeros((len(seqs), EO.MAX_LEN), np.uint16)
    L = OB.ref()
    L.ref_tokenize.restype = C.c_int
    rc = L.ref_tokenize(C.c_char_p(flat), off.ctypes.data_as(C.c_void_p), C.c_int64(len(seqs)), C.c_int(EO.MAX_LEN),
                        ref_tok.ctypes.data_as(C.c_void_p))
    assert rc == 0
    mine = EO.tokenise_batch(seqs)
    assert np.array_equal(mine, ref_tok), "tokeniser restatement differs from the reference"
    used = np.unique(ref_tok)
    print(f"tokeniser == reference on {len(seqs)} sequences; {len(used)} distinct ids")

    # 3. the network vs torch.nn.GRU
    import torch
    torch.set_num_threads(4)
    H = 64

    def zrh_to_rzn(a):  # rows [z, r, h] -> torch's [r, z, n]
        return np.concatenate([a[H:2 * H], a[0:H], a[2 * H:3 * H]], axis=0)

    gru = torch.nn.GRU(64, 64, num_layers=2, bidirectional=True)
    sd = {}
    for layer, (Wk, Rk, Bk) in enumerate((("W1", "R1", "B1"), ("W2", "R2", "B2"))):
        for d, suf in enumerate(("", "_reverse")):
            W = w[Wk][d].astype(np.float32); R = w[Rk][d].astype(np.float32); B = w[Bk][d].astype(np.float32)
            sd[f"weight_ih_l{layer}{suf}"] = torch.from_numpy(zrh_to_rzn(W))
            sd[f"weight_hh_l{layer}{suf}"] = torch.from_numpy(zrh_to_rzn(R))
            # folded z/r biases go to b_ih, b_hh gets zeros there; the h gate keeps both
            b_ih = np.concatenate([B[H:2 * H], B[0:H], B[2 * H:3 * H]])
            b_hh = np.concatenate([np.zeros(2 * H, np.float32), B[3 * H:4 * H]])
            sd[f"bias_ih_l{layer}{suf}"] = torch.from_numpy(b_ih)
            sd[f"bias_hh_l{layer}{suf}"] = torch.from_numpy(b_hh)
    gru.load_state_dict(sd)
    pick = [i for i, s in enumerate(seqs) if len(s) >= 100][:48] + list(range(0, 80, 5))
    tok = ref_tok[pick].astype(np.int64)
    with torch.no_grad():
        x = torch.from_numpy(w["emb"].astype(np.float32))[torch.from_numpy(tok)].transpose(0, 1)  # [T, n, 64]
        _, hn = gru(x)
        want = torch.cat([hn[2], hn[3]], dim=1).numpy()
    got = EO.embed_tokens(tok, w)
    err = float(np.abs(got - want).max())
    print(f"numpy restatement vs torch.nn.GRU: max abs diff {err:.3e} over {got.shape}")
    assert err < 2e-5

    ids = np.array([0] + [EO.token_id(s) for s in range(96)], np.int32)
    assert set(used.tolist()) <= set(ids.tolist())
    out = os.path.join(ROOT, "tests", "golden", "embedder_golden.npz")
    np.savez_compressed(
        out, emb_ids=ids, emb_rows=w["emb"][ids], W1=w["W1"], R1=w["R1"], B1=w["B1"], W2=w["W2"], R2=w["R2"], B2=w["B2"],
        seq_flat=np.frombuffer(b"".join(seqs[i] for i in pick), np.uint8),
        seq_off=np.concatenate([[0], np.cumsum([len(seqs[i]) for i in pick])]).astype(np.int64),
        tokens=ref_tok[pick], embeddings=want.astype(np.float32),
        tok_seq_flat=np.frombuffer(flat, np.uint8), tok_seq_off=off, tok_all=ref_tok)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
