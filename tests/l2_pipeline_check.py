"""The C++ host code of `pipeline`'s DRM_RERANK=l2 stage (post_process_l2_flat in host/stage.cpp, through
`formats_tool l2flat`) against oracle/l2_oracle.py: dynamic over a FASTA genome, static over the bare window table
read_file(..., lookup_mode=true) builds from the same FASTA (device mode 2).  Run as a script: on a GPU box by tests/test_gpu_zz_l2.py (child process), here by
tests/test_emu_kernels.py with LD_PRELOAD = the host-thread emulation of the same library.

Candidate embeddings on both sides come from the library's own embedder (validated separately), so ids and fp32
distance bits must be equal."""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import deepreadmapper_b200 as drm  # noqa: E402
import embed_util as U  # noqa: E402
from oracle import embed_oracle as EO  # noqa: E402
from oracle import l2_oracle as L2  # noqa: E402


def bits(a):
    return a.view(np.uint32) if a.dtype == np.float32 else a


def main(n=6, preload=None):
    g = np.random.default_rng(71)
    tmp = tempfile.mkdtemp(prefix="drm_l2flat_")
    w = EO.random_weights(72)
    xml = U.write_ir(os.path.join(tmp, "model"), w)
    v = drm.Vectorizer(model_path=xml)
    embed_fn = lambda seqs: v.vectorize(seqs)  # noqa: E731
    ref_len, k_in = 150, 6
    genome = bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), 1500))
    fasta = os.path.join(tmp, "ref.fna")
    with open(fasta, "wb") as f:
        f.write(b"> synthetic\n" + b"\n".join(genome[i:i + 70] for i in range(0, len(genome), 70)) + b"\n")
    q_emb = embed_fn([b"<" + genome[40 * i:40 * i + 150] + b">" for i in range(n)])
    dist = np.sort(g.random((n, k_in)).astype(np.float32), axis=1)
    np.save(os.path.join(tmp, "D.npy"), dist)
    np.save(os.path.join(tmp, "Q.npy"), q_emb)
    env = dict(os.environ, DRM_MODEL=xml)
    if preload:
        env["LD_PRELOAD"] = preload
    tool = os.path.join(ROOT, "build", "bin", "formats_tool")

    def run(mode, nb, stride, k, k_clusters):
        np.save(os.path.join(tmp, "I.npy"), nb.astype(np.int64))
        r = subprocess.run([tool, "l2flat", str(mode), fasta, str(ref_len), os.path.join(tmp, "I.npy"), os.path.join(tmp, "D.npy"),
                            os.path.join(tmp, "Q.npy"), str(stride), str(k), str(k_clusters), os.path.join(tmp, "s.npy"), os.path.join(tmp, "i.npy")],
                           capture_output=True, text=True, env=env, timeout=1200)
        assert r.returncode == 0, r.stderr[-2000:] + r.stdout[-500:]
        return np.load(os.path.join(tmp, "s.npy")), np.load(os.path.join(tmp, "i.npy"))

    # dynamic (mode 1), sparse index
    stride, k_clusters, k = 4, 4, 9
    nb = g.integers(1, (len(genome) - stride) // stride, (n, k_in)).astype(np.int64)
    s, i = run(1, nb, stride, k, k_clusters)
    s0, i0 = L2.post_process_l2_dynamic(nb, dist, genome, ref_len, stride, k, k_clusters, q_emb, embed_fn)
    assert i.dtype == np.uint64 and s.dtype == np.float32 and s.ndim == 1
    assert np.array_equal(i.astype(np.int64), np.asarray(i0).astype(np.int64)) and np.array_equal(bits(s), bits(np.asarray(s0, np.float32))), "dynamic"
    print("l2flat dynamic ok:", len(i), "values")

    # static (mode 0): the window table is what read_file(fasta, ref_len, 1, lookup) returns
    comp = bytes.maketrans(b"ACGT", b"TGCA")
    ref_seqs = []
    for p in range(len(genome) - ref_len + 1):
        win = genome[p:p + ref_len]
        ref_seqs += [win, win.translate(comp)[::-1]]  # lookup_mode windows carry no '<' '>' (src/utils/parse_inputs.cpp:345-349)
    stride, kk = 3, 4
    nb = g.integers(1, (len(ref_seqs) - stride) // stride, (n, k_in)).astype(np.int64)
    s, i = run(0, nb, stride, kk, kk)
    s0, i0 = L2.post_process_l2_static(nb, dist, ref_seqs, ref_len, stride, kk, kk, q_emb, embed_fn)
    assert np.array_equal(i.astype(np.int64), np.asarray(i0).astype(np.int64)) and np.array_equal(bits(s), bits(np.asarray(s0, np.float32))), "static"
    print("l2flat static ok:", len(i), "values")
    v.close()
    print("pipeline L2 stage: host code == oracle")


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 6, sys.argv[2] if len(sys.argv) > 2 else None)
