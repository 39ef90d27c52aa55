"""First hardware check of the L2 rerank stage (drm_l2_rerank) against oracle/l2_oracle.py.  Run as a script by
tests/test_gpu_zz_l2.py in a child process, so that a device fault cannot take the test session's CUDA context with it.

The embeddings on both sides come from the (already validated) device embedder, so everything after them - candidate
bookkeeping, undecorated / decorated window fetch by id, the unfused distance chain, partial_sort's order - must agree
bit for bit: ids and fp32 distance bits."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import deepreadmapper_b200 as drm  # noqa: E402
from oracle import embed_oracle as EO  # noqa: E402
from oracle import l2_oracle as L2  # noqa: E402


def same(a, b):
    return a.shape == b.shape and np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a, b.view(np.uint32) if b.dtype == np.float32 else b)


def main(n=23):
    g = np.random.default_rng(61)
    v = drm.Vectorizer(weights=EO.random_weights(62))
    embed_fn = lambda seqs: v.vectorize(seqs)  # noqa: E731
    ref_len = 150
    # a genome with a repeated 400-base segment: identical windows at different positions -> equal distances, different ids.
    # The dynamic variant compares id * stride with the genome LENGTH (src/utils/post_processor.cpp:117-121), so only dense
    # ids below it - positions in the first half - are reachable; both copies sit there.
    seg = bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), 400))
    genome = bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), 300)) + seg + bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), 500)) + seg + \
        bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), 2400))
    ref = drm.Reference(genome=genome, ref_len=ref_len)
    k_in = 7
    q_emb = embed_fn([b"<" + genome[90 * i:90 * i + 150] + b">" for i in range(n)])
    dist = np.sort(g.random((n, k_in)).astype(np.float32), axis=1)

    # dynamic, sparse index: labels whose expansions are complete; some land in both copies of the repeat
    stride, k_clusters, k = 4, 5, 12
    nb = g.integers(1, (len(genome) - stride) // stride, (n, k_in)).astype(np.int64)
    nb[:, 0] = (2 * 350 + np.arange(n) * 8) // stride  # windows inside the first copy of the repeat (dense id = 2 * position + strand)
    nb[:, 1] = (2 * 1250 + np.arange(n) * 8) // stride  # the same windows in the second copy (300 + 400 + 500 + 50)
    s, i = drm.post_process_l2_dynamic(nb, dist, ref, q_emb, v, stride, k, k_clusters)
    s0, i0 = L2.post_process_l2_dynamic(nb, dist, genome, ref_len, stride, k, k_clusters, q_emb, embed_fn)
    assert same(i, i0) and same(s, s0), "dynamic sparse"
    ties = sum(int(np.any(np.diff(s0[q * k:(q + 1) * k]) == 0)) for q in range(n))
    print("dynamic sparse ok; queries with equal distances in the top-k:", ties)

    # dynamic, dense index: pass-through
    s, i = drm.post_process_l2_dynamic(nb, dist, ref, q_emb, v, 1, 5, 5)
    s0, i0 = L2.post_process_l2_dynamic(nb, dist, genome, ref_len, 1, 5, 5, q_emb, embed_fn)
    assert same(i, i0) and same(s, s0), "dynamic dense"

    # not enough candidates: k above what one label expands to
    try:
        drm.post_process_l2_dynamic(nb[:, :1], dist[:, :1], ref, q_emb, v, stride, 8, 1)
        raise AssertionError("expected DRM_ENOTENOUGH")
    except drm.DrmError as e:
        assert e.code == -5 and "Not enough candidates (7 < 8) for query 0" in str(e), str(e)

    # static on the dense windows of the same genome (ref_seqs = '<' window '>' strings, forward / reverse per position)
    comp = bytes.maketrans(b"ACGT", b"TGCA")
    ref_seqs = []
    for p in range(len(genome) - ref_len + 1):
        w = genome[p:p + ref_len]
        ref_seqs += [b"<" + w + b">", b"<" + w.translate(comp)[::-1] + b">"]
    nb = g.integers(0, len(ref_seqs), (n, k_in)).astype(np.int64)
    for stride, kk in ((1, 5), (3, 4)):
        if stride > 1:
            nb = g.integers(1, (len(ref_seqs) - stride) // stride, (n, k_in)).astype(np.int64)
        s, i = drm.post_process_l2_static(nb, dist, ref, q_emb, v, stride, kk, kk)
        s0, i0 = L2.post_process_l2_static(nb, dist, ref_seqs, ref_len, stride, kk, kk, q_emb, embed_fn)
        assert same(i, i0) and same(s, s0), f"static stride {stride}"
    # ... and on an explicit window table
    table = drm.Reference(windows=[x[1:-1] for x in ref_seqs[:900]], ref_len=ref_len)  # noqa: E501
    nb = g.integers(0, 900, (n, k_in)).astype(np.int64)
    s, i = drm.post_process_l2_static(nb, dist, table, q_emb, v, 1, 5, 5)
    s0, i0 = L2.post_process_l2_static(nb, dist, ref_seqs[:900], ref_len, 1, 5, 5, q_emb, embed_fn)
    assert same(i, i0) and same(s, s0), "static table"
    # ... and the same table the way the reference's main loads it (read_file in lookup mode: bare windows, embedded undecorated)
    bare_seqs = [x[1:-1] for x in ref_seqs[:900]]
    s, i = drm.post_process_l2_static(nb, dist, table, q_emb, v, 1, 5, 5, decorated=False)
    s0, i0 = L2.post_process_l2_static(nb, dist, bare_seqs, ref_len, 1, 5, 5, q_emb, embed_fn)
    assert same(i, i0) and same(s, s0), "static table, bare windows (mode 2)"
    sd, _ = drm.post_process_l2_static(nb, dist, table, q_emb, v, 1, 5, 5)
    assert not same(s, sd), "decoration must change the embeddings"
    print("L2 rerank stage: device == oracle")


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 23)
