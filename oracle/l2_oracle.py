"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference's L2 rerank stage (the post-process variants that are
live in src/main.cpp and need the embedder).  Never imported by the product.

    post_process_l2_dynamic   src/utils/post_processor.cpp:552-748   (find_sequences dynamic :74-201)
    post_process_l2_static    src/utils/post_processor.cpp:1023-1162 (find_sequences static  :204-336)
    batch_reranker            src/utils/reranker.cpp:98-196          (numeric core: oracle/l2_oracle.c)

The restatement follows the reference's FLAT bookkeeping literally, because that is what decides which candidates a
query sees: all queries' sparse ids are expanded into one list (out-of-range ids are skipped, windows at the start of
the reference are clipped), while the per-query boundaries are computed from a fixed count per id -
(2*stride - 1) in the dynamic variant, `stride` in the static one.  When every id expands fully the dynamic variant's
boundaries are exact; otherwise (and always for static with stride > 1) a query's slice is shifted.  Slices that would
run past the flat list are undefined behaviour in the reference; here they raise.

Pinned against the reference's own code: oracle/_ref compiles post_processor.cpp / reranker.cpp / metrics.cpp unmodified,
with Vectorizer::vectorize routed to a callback (oracle/ref_shim.cpp: ref_set_vectorizer) that the tests point at
oracle/embed_oracle.py - so everything except the network itself is the reference (tests/test_oracle_l2.py).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import bindings as B

U64 = (1 << 64) - 1
_COMP = bytes.maketrans(b"ATCGN", b"TAGCN")


class OracleError(RuntimeError):
    pass


def _lib():
    L = B.oracle()
    if not getattr(L, "_l2_ready", False):
        L.orc_l2_dist.restype = C.c_float
        L.orc_l2_dist.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int]
        L.orc_partial_sort_asc.restype = None
        L.orc_partial_sort_asc.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
        L.orc_batch_rerank_one.restype = C.c_int64
        L.orc_batch_rerank_one.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64,
                                           C.c_int, C.c_void_p, C.c_void_p]
        L._l2_ready = True
    return L


def l2_dist(a, b, use_fma=False) -> float:
    a = np.ascontiguousarray(a, np.float32); b = np.ascontiguousarray(b, np.float32)
    return float(_lib().orc_l2_dist(a.ctypes.data, b.ctypes.data, a.size, int(use_fma)))


def partial_sort_asc(dist, k):
    dist = np.ascontiguousarray(dist, np.float32)
    idx = np.zeros(len(dist), np.int64)
    _lib().orc_partial_sort_asc(dist.ctypes.data, len(dist), k, idx.ctypes.data)
    return idx[:k]


def find_sequence(genome: bytes, dense_id: int, ref_len: int) -> bytes:
    """src/utils/post_processor.cpp:47-66 (bytes outside ATCGN complement to NUL, src/utils/parse_inputs.cpp:5-14)."""
    pos = dense_id // 2
    if pos + ref_len > len(genome):
        return b""
    w = genome[pos:pos + ref_len]
    if dense_id % 2:
        w = bytes(c if chr(c) in "ATCGN" else 0 for c in w.translate(_COMP))[::-1]
    return w


def _expand_flat(sparse_ids, stride: int, bound: int):
    """The sparse branch of find_sequences: (unique dense ids in order of first appearance, flat mapping into them)."""
    order, index, mapping = [], {}, []
    for sid in sparse_ids:
        p = (int(sid) * stride) & U64  # size_t arithmetic
        if p >= bound:
            continue
        start = p - stride + 1 if p >= stride - 1 else 0
        end = min(p + stride, bound)
        for pos in range(start, end):
            if pos not in index:
                index[pos] = len(order)
                order.append(pos)
            mapping.append(index[pos])
    return order, mapping


def _rerank_slices(unique_emb, dense_ids, mapping, starts, q_emb, k, use_fma):
    L = _lib()
    unique_emb = np.ascontiguousarray(unique_emb, np.float32)
    mapping = np.asarray(mapping, np.uint64)
    dense_of_flat = np.asarray(dense_ids, np.uint64)[mapping.astype(np.int64)] if len(mapping) else np.zeros(0, np.uint64)
    out_d, out_i = [], []
    for q in range(len(q_emb)):
        a, b = starts[q], starts[q + 1]
        if b > len(mapping):
            raise OracleError(f"query {q}: slice [{a}, {b}) runs past the {len(mapping)} flat candidates (undefined in the reference)")
        n_c = b - a
        if n_c == 0:
            continue
        d = np.zeros(k, np.float32); i = np.zeros(k, np.uint64)
        m = np.ascontiguousarray(mapping[a:b]); di = np.ascontiguousarray(dense_of_flat[a:b])
        qe = np.ascontiguousarray(q_emb[q], np.float32)
        rc = L.orc_batch_rerank_one(unique_emb.ctypes.data, m.ctypes.data, di.ctypes.data, n_c, qe.ctypes.data, unique_emb.shape[1], k,
                                    int(use_fma), d.ctypes.data, i.ctypes.data)
        if rc < 0:
            raise OracleError(f"Not enough candidates ({n_c} < {k}) for query {q}")
        out_d.append(d); out_i.append(i)
    if not out_d:
        return np.zeros(0, np.float32), np.zeros(0, np.uint64)
    return np.concatenate(out_d), np.concatenate(out_i)


def post_process_l2_dynamic(neighbors, distances, genome: bytes, ref_len, stride, k, k_clusters, q_emb, embed_fn, use_fma=False):
    """Returns (scores float32, dense ids uint64), flat like the reference's final vectors.  embed_fn(list of bytes) ->
    [n, 128] float32 stands for Vectorizer::vectorize; it receives the UNDECORATED windows find_sequence returns."""
    neighbors = np.asarray(neighbors, np.int64)
    n, k_in = neighbors.shape
    if k > k_clusters * 2 * stride and stride > 1:
        raise OracleError("Final k too large. Ensure k < k_clusters * 2 * stride to have enough candidates.")
    per_q = min(k if stride == 1 else k_clusters, k_in)
    sparse = neighbors[:, :per_q].astype(np.uint64).reshape(-1)  # long -> size_t
    if stride == 1:  # "stride == 1 (dense), skipping reranker": the search order and distances pass through
        take = min(k, per_q)
        return (np.asarray(distances, np.float32)[:, :take].reshape(-1).copy(),
                neighbors[:, :take].astype(np.uint64).reshape(-1).copy())
    order, mapping = _expand_flat(sparse, stride, len(genome))
    emb = embed_fn([find_sequence(genome, p, ref_len) for p in order]) if order else np.zeros((0, 128), np.float32)
    starts = [q * per_q * (2 * stride - 1) for q in range(n + 1)]
    return _rerank_slices(emb, order, mapping, starts, q_emb, k, use_fma)


def post_process_l2_static(neighbors, distances, ref_seqs, ref_len, stride, k, k_clusters, q_emb, embed_fn, use_fma=False):
    """ref_seqs: list of bytes as read_file returns them (decorated).  Returns k_clusters results per query - the static
    variant hands k_clusters, not k, to batch_reranker (src/utils/post_processor.cpp:1128-1135)."""
    neighbors = np.asarray(neighbors, np.int64)
    n, k_in = neighbors.shape
    ids = neighbors.astype(np.uint64).reshape(-1)
    n_ref = len(ref_seqs)
    if stride == 1:
        keep = [j for j, v in enumerate(ids) if int(v) < n_ref]
        order = [int(ids[j]) for j in keep]
        mapping = keep  # the reference stores the ORIGINAL position, not the compacted one (:232)
        if any(m >= len(order) for m in mapping):
            raise OracleError("Invalid mapping index in expansion")
    else:
        order, mapping = _expand_flat(ids, stride, n_ref)
    emb = embed_fn([ref_seqs[p] for p in order]) if order else np.zeros((0, 128), np.float32)
    per = k_in * (stride if stride > 1 else 1)
    starts = [q * per for q in range(n + 1)]
    return _rerank_slices(emb, order, mapping, starts, q_emb, k_clusters, use_fma)
