"""Helpers for the embedder tests: the golden fixture, and a writer for a minimal OpenVINO IR pair of the embedder's
architecture (same layer types, ports and <data> attributes the reference's models/*.xml uses for the tensors the
loader reads; the shape-juggling layers in between are left out)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "embedder_golden.npz")
VOCAB = 7638


def load_golden():
    d = np.load(GOLDEN)
    w = {k: d[k] for k in ("W1", "R1", "B1", "W2", "R2", "B2")}
    emb = np.zeros((VOCAB, 64), np.float16)  # only the rows the tokeniser can emit are in the fixture
    emb[d["emb_ids"]] = d["emb_rows"]
    w["emb"] = emb
    off = d["seq_off"]
    flat = d["seq_flat"].tobytes()
    seqs = [flat[off[i]:off[i + 1]] for i in range(len(off) - 1)]
    toff = d["tok_seq_off"]
    tflat = d["tok_seq_flat"].tobytes()
    tok_seqs = [tflat[toff[i]:toff[i + 1]] for i in range(len(toff) - 1)]
    return {"weights": w, "seqs": seqs, "tokens": d["tokens"], "embeddings": d["embeddings"], "tok_seqs": tok_seqs,
            "tok_all": d["tok_all"]}


def write_ir(path_base, w, f32=()):
    """Writes path_base.xml / .bin.  Tensors named in `f32` are stored as f32 constants (no Convert), the rest as f16
    constants behind a Convert, as the reference's compressed IR has them."""
    blob = bytearray()
    layers, edges = [], []
    next_id = [0]

    def new_id():
        next_id[0] += 1
        return next_id[0] - 1

    def const(name, arr):
        as32 = name in f32
        a = np.ascontiguousarray(arr, dtype=np.float32 if as32 else np.float16)
        while len(blob) % 8:  # the real file has gaps between tensors too
            blob.append(0)
        off = len(blob)
        blob.extend(a.tobytes())
        cid = new_id()
        shape = ", ".join(str(x) for x in a.shape)
        layers.append(f'<layer id="{cid}" name="{name}_c" type="Const" version="opset1">\n'
                      f'\t<data element_type="{"f32" if as32 else "f16"}" shape="{shape}" offset="{off}" size="{a.nbytes}" />\n'
                      f'\t<output><port id="0" precision="FP16" /></output>\n</layer>')
        if as32:
            return cid
        vid = new_id()
        layers.append(f'<layer id="{vid}" name="{name}" type="Convert" version="opset1">\n'
                      f'\t<data destination_type="f32" />\n</layer>')
        edges.append((cid, 0, vid, 0))
        return vid

    pid = new_id()
    layers.append(f'<layer id="{pid}" name="input" type="Parameter" version="opset1">\n'
                  f'\t<data shape="123,100" element_type="i64" />\n</layer>')
    emb = const("emb.weight", w["emb"])
    gid = new_id()
    layers.append(f'<layer id="{gid}" name="/emb/Gather" type="Gather" version="opset8">\n\t<data batch_dims="0" />\n</layer>')
    edges += [(emb, 1 if "emb.weight" not in f32 else 0, gid, 0), (pid, 0, gid, 1)]
    prev = gid
    for layer, (Wk, Rk, Bk) in enumerate((("W1", "R1", "B1"), ("W2", "R2", "B2"))):
        W, R, B = const(Wk, w[Wk]), const(Rk, w[Rk]), const(Bk, w[Bk])
        rid = new_id()
        layers.append(f'<layer id="{rid}" name="GRUSequence_{layer}" type="GRUSequence" version="opset5">\n'
                      f'\t<data direction="bidirectional" linear_before_reset="true" hidden_size="64" '
                      f'activations="sigmoid, tanh" activations_alpha="" activations_beta="" clip="0" />\n</layer>')
        edges += [(prev, 0, rid, 0), (W, 1, rid, 3), (R, 1, rid, 4), (B, 1, rid, 5)]
        prev = rid
    out = new_id()
    layers.append(f'<layer id="{out}" name="output/sink" type="Result" version="opset1" />')
    edges.append((prev, 6, out, 0))
    xml = ['<?xml version="1.0"?>', '<net name="embedder" version="11">', "<layers>"] + layers + ["</layers>", "<edges>"]
    xml += [f'<edge from-layer="{a}" from-port="{b}" to-layer="{c}" to-port="{d}" />' for a, b, c, d in edges]
    xml += ["</edges>", "</net>"]
    with open(path_base + ".xml", "w") as f:
        f.write("\n".join(xml) + "\n")
    with open(path_base + ".bin", "wb") as f:
        f.write(bytes(blob))
    return path_base + ".xml"
