"""CPU tests: oracle/l2_oracle.{c,py} - the restatement of the reference's L2 rerank stage - against the reference's own
post_process_l2_static / post_process_l2_dynamic / batch_reranker / calc_l2_dist compiled into oracle/_ref, with
Vectorizer::vectorize routed to oracle/embed_oracle.py (the one piece OpenVINO would supply).  This is the oracle for the
next row to move onto the GPU; there is no CUDA path for it yet."""
import ctypes as C

import numpy as np
import pytest

from oracle import bindings as OB
from oracle import embed_oracle as EO
from oracle import l2_oracle as L2

pytestmark = pytest.mark.skipif(not OB.have_ref(), reason="oracle/_ref is not built (needs /root/reference once)")

CB = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(C.c_int64), C.c_int64, C.POINTER(C.c_float))


@pytest.fixture(scope="module")
def env():
    w = EO.random_weights(7)
    calls = []

    def embed_fn(seqs):
        calls.append(len(seqs))
        return EO.embed(seqs, w)

    def cb(flat, off, n, out):
        offs = [off[i] for i in range(n + 1)]
        raw = C.string_at(flat, offs[-1]) if offs[-1] else b""
        e = embed_fn([raw[offs[i]:offs[i + 1]] for i in range(n)])
        C.memmove(out, e.ctypes.data, e.nbytes)

    keep = CB(cb)
    R = OB.ref()
    R.ref_set_vectorizer.argtypes = [CB]
    R.ref_set_vectorizer.restype = None
    R.ref_set_vectorizer(keep)
    R.ref_calc_l2_dist.restype = C.c_float
    R.ref_calc_l2_dist.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    R.ref_post_process_l2.restype = C.c_int
    R.ref_post_process_l2.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_char_p, C.c_int64, C.c_int64, C.c_int64,
                                      C.c_char_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                      C.POINTER(C.c_int64)]
    yield {"R": R, "embed_fn": embed_fn, "keep": keep, "w": w}
    R.ref_set_vectorizer(C.cast(None, CB))


def _ref_l2(R, mode, nb, dist, ref_flat, n_ref, ref_w, ref_len, q_emb, stride, k, k_clusters):
    n, k_in = nb.shape
    q_flat = b"<ACGT>" * n  # query strings are carried but not used by the L2 variants
    q_off = np.arange(n + 1, dtype=np.int64) * 6
    cap = n * max(k, k_clusters) + 8
    scores = np.zeros(cap, np.float32); ids = np.zeros(cap, np.uint64); cnt = C.c_int64(0)
    nb = np.ascontiguousarray(nb, np.int64); dist = np.ascontiguousarray(dist, np.float32); q_emb = np.ascontiguousarray(q_emb, np.float32)
    rc = R.ref_post_process_l2(mode, nb.ctypes.data, dist.ctypes.data, n, k_in, ref_flat, n_ref, ref_w, ref_len, q_flat, q_off.ctypes.data,
                               q_emb.ctypes.data, stride, k, k_clusters, scores.ctypes.data, ids.ctypes.data, C.byref(cnt))
    if rc != 0:
        return rc, R.ref_last_error().decode(), None
    return 0, scores[:cnt.value].copy(), ids[:cnt.value].copy()


def test_l2_distance_is_the_unfused_ordered_chain(env):
    g = np.random.default_rng(1)
    A = g.standard_normal((300, 128)).astype(np.float32); Bm = g.standard_normal((300, 128)).astype(np.float32)
    ref = np.array([env["R"].ref_calc_l2_dist(A[i].ctypes.data, Bm[i].ctypes.data, 128) for i in range(300)], np.float32)
    fused = np.array([L2.l2_dist(A[i], Bm[i], True) for i in range(300)], np.float32)
    plain = np.array([L2.l2_dist(A[i], Bm[i], False) for i in range(300)], np.float32)
    # g++ vectorises the subtract / multiply and adds in order with vaddss: products are rounded first, no FMA
    assert np.array_equal(ref.view(np.uint32), plain.view(np.uint32))
    assert not np.array_equal(fused.view(np.uint32), plain.view(np.uint32))
    assert np.allclose(plain, np.sqrt(((A.astype(np.float64) - Bm) ** 2).sum(1)), rtol=1e-6)


def test_partial_sort_asc_orders_and_breaks_ties_like_heap_select():
    g = np.random.default_rng(2)
    for n, k in ((1, 1), (7, 3), (35, 10), (35, 35), (200, 16)):
        d = g.integers(0, 6, n).astype(np.float32)  # many ties
        idx = L2.partial_sort_asc(d, k)
        assert len(set(idx.tolist())) == k
        assert np.array_equal(d[idx], np.sort(d)[:k])


def test_dynamic_variant_against_reference(env):
    g = np.random.default_rng(3)
    genome = bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), 3000))
    ref_len, stride, k_clusters, k, n, k_in = 150, 4, 5, 12, 9, 7
    # sparse labels whose windows expand fully: stride-1 <= id*stride and id*stride + stride <= len(genome)
    nb = g.integers(1, (len(genome) - stride) // stride, (n, k_in)).astype(np.int64)
    dist = np.sort(g.random((n, k_in)).astype(np.float32), axis=1)
    q_emb = env["embed_fn"]([b"<" + genome[100 * i:100 * i + 150] + b">" for i in range(n)])
    rc, s_ref, i_ref = _ref_l2(env["R"], 1, nb, dist, genome, len(genome), 0, ref_len, q_emb, stride, k, k_clusters)
    assert rc == 0, s_ref
    s, i = L2.post_process_l2_dynamic(nb, dist, genome, ref_len, stride, k, k_clusters, q_emb, env["embed_fn"])
    assert len(s) == n * k and np.array_equal(i, i_ref) and np.array_equal(s.view(np.uint32), s_ref.view(np.uint32))
    # every returned id is one of the query's expansions, distances ascend
    for q in range(n):
        exp = set()
        for sid in nb[q, :k_clusters]:
            exp.update(range(int(sid) * stride - stride + 1, int(sid) * stride + stride))
        assert set(i[q * k:(q + 1) * k].tolist()) <= exp
        assert np.all(np.diff(s[q * k:(q + 1) * k]) >= 0)
    # dense index: the reranker is skipped, search order and distances pass through (:638-664)
    rc, s_ref, i_ref = _ref_l2(env["R"], 1, nb, dist, genome, len(genome), 0, ref_len, q_emb, 1, 5, 5)
    s, i = L2.post_process_l2_dynamic(nb, dist, genome, ref_len, 1, 5, 5, q_emb, env["embed_fn"])
    assert rc == 0 and np.array_equal(i, i_ref) and np.array_equal(s, s_ref) and np.array_equal(i.reshape(n, 5), nb[:, :5].astype(np.uint64))
    # the guard
    rc, msg, _ = _ref_l2(env["R"], 1, nb, dist, genome, len(genome), 0, ref_len, q_emb, stride, 41, 5)
    assert rc != 0 and msg.startswith("Final k too large")
    with pytest.raises(L2.OracleError, match="Final k too large"):
        L2.post_process_l2_dynamic(nb, dist, genome, ref_len, stride, 41, 5, q_emb, env["embed_fn"])


def test_static_variant_against_reference(env):
    g = np.random.default_rng(4)
    genome = bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), 700))
    ref_len = 150
    comp = bytes.maketrans(b"ACGT", b"TGCA")
    ref_seqs = []
    for p in range(len(genome) - ref_len + 1):
        w = genome[p:p + ref_len]
        ref_seqs += [b"<" + w + b">", b"<" + w.translate(comp)[::-1] + b">"]
    flat = b"".join(ref_seqs)
    n, k_in = 6, 8
    q_emb = env["embed_fn"]([ref_seqs[37 * i] for i in range(n)])
    # dense: a real rerank of the k_in neighbours by embedding distance, k_clusters results per query
    nb = g.integers(0, len(ref_seqs), (n, k_in)).astype(np.int64)
    dist = np.sort(g.random((n, k_in)).astype(np.float32), axis=1)
    rc, s_ref, i_ref = _ref_l2(env["R"], 0, nb, dist, flat, len(ref_seqs), ref_len + 2, ref_len, q_emb, 1, 5, 5)
    assert rc == 0, s_ref
    s, i = L2.post_process_l2_static(nb, dist, ref_seqs, ref_len, 1, 5, 5, q_emb, env["embed_fn"])
    assert len(s) == n * 5 and np.array_equal(i, i_ref) and np.array_equal(s.view(np.uint32), s_ref.view(np.uint32))
    # sparse: the reference's boundaries count `stride` candidates per id while every id expands to 2*stride-1, so a
    # query's slice is shifted - restated as is
    stride = 3
    nb = g.integers(1, (len(ref_seqs) - stride) // stride, (n, k_in)).astype(np.int64)
    rc, s_ref, i_ref = _ref_l2(env["R"], 0, nb, dist, flat, len(ref_seqs), ref_len + 2, ref_len, q_emb, stride, 4, 4)
    assert rc == 0, s_ref
    s, i = L2.post_process_l2_static(nb, dist, ref_seqs, ref_len, stride, 4, 4, q_emb, env["embed_fn"])
    assert len(s) == n * 4 and np.array_equal(i, i_ref) and np.array_equal(s.view(np.uint32), s_ref.view(np.uint32))


def test_undefined_slices_raise_in_the_oracle(env):
    g = np.random.default_rng(5)
    genome = bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), 2000))
    nb = g.integers(1, 400, (4, 6)).astype(np.int64)
    nb[0, 0] = 0  # clipped at the start of the genome: fewer expansions than the boundaries assume
    q_emb = np.zeros((4, 128), np.float32)
    with pytest.raises(L2.OracleError, match="runs past"):
        L2.post_process_l2_dynamic(nb, np.zeros((4, 6), np.float32), genome, 150, 4, 8, 5, q_emb, env["embed_fn"])


# ---- the product's host-side pieces of the L2 stage (no device needed) against the oracle ---------------------------------
def test_product_selection_order_equals_oracle_heap_select():
    import deepreadmapper_b200 as drm
    g = np.random.default_rng(6)
    for n, k in ((1, 1), (2, 2), (7, 3), (35, 10), (35, 35), (75, 12), (480, 128), (300, 1)):
        for ties in (True, False):
            d = (g.integers(0, 5, n) if ties else g.random(n)).astype(np.float32)
            assert np.array_equal(drm.l2_select_host(d, k), L2.partial_sort_asc(d, k)), (n, k, ties)


@pytest.mark.parametrize("table", ["direct", "hash"])
def test_product_plan_equals_oracle_bookkeeping(table, monkeypatch):
    """Both position tables of l2_plan (direct by position - the default below 2^29 positions - and open addressing)."""
    import deepreadmapper_b200 as drm
    monkeypatch.setenv("DRM_L2_PLAN_TABLE", table)
    g = np.random.default_rng(7)
    n, k_in = 11, 9
    # dynamic, sparse
    stride, k_clusters, k, bound = 4, 5, 12, 3000
    nb = g.integers(1, (bound - stride) // stride, (n, k_in)).astype(np.int64)
    p = drm.l2_plan_host(1, nb, stride, k, k_clusters, bound)
    order, mapping = L2._expand_flat(nb[:, :k_clusters].astype(np.uint64).reshape(-1), stride, bound)
    assert p["k_out"] == k and not p["passthrough"]
    assert p["unique"].tolist() == order and p["map"].tolist() == mapping
    assert p["starts"].tolist() == [q * k_clusters * (2 * stride - 1) for q in range(n + 1)]
    # dynamic, dense: pass-through of min(k, k_in) search results
    p = drm.l2_plan_host(1, nb, 1, 5, 5, bound)
    assert p["passthrough"] and p["take"] == 5 and p["k_out"] == 5
    # static, dense (drops nothing here) and sparse (boundaries count `stride` per id)
    p = drm.l2_plan_host(0, nb, 1, 5, 5, bound)
    assert p["k_out"] == 5 and p["unique"].tolist() == nb.reshape(-1).tolist() and p["map"].tolist() == list(range(n * k_in))
    assert p["starts"].tolist() == [q * k_in for q in range(n + 1)]
    p = drm.l2_plan_host(0, nb, 3, 4, 4, bound)
    order, mapping = L2._expand_flat(nb.astype(np.uint64).reshape(-1), 3, bound)
    assert p["unique"].tolist() == order and p["map"].tolist() == mapping and p["starts"].tolist() == [q * k_in * 3 for q in range(n + 1)]
    # the plan made when every window's embedding is cached: the flat list carries the window ids, same order, same boundaries
    for mode, st, kk, kc in ((1, stride, k, k_clusters), (0, 3, 4, 4)):
        a = drm.l2_plan_host(mode, nb, st, kk, kc, bound)
        b = drm.l2_plan_host(mode, nb, st, kk, kc, bound, direct_ids=True)
        assert len(b["unique"]) == 0 and b["map"].tolist() == a["unique"][a["map"]].tolist() and b["starts"].tolist() == a["starts"].tolist()
    # errors: the guard, and slices that would leave the flat list (a clipped id)
    with pytest.raises(drm.DrmError, match="Final k too large") as e:
        drm.l2_plan_host(1, nb, stride, 41, 5, bound)
    assert e.value.code == -6
    nb2 = nb.copy(); nb2[0, 0] = 0
    with pytest.raises(drm.DrmError, match="run past") as e:
        drm.l2_plan_host(1, nb2, stride, k, k_clusters, bound)
    assert e.value.code == -1
    nb3 = nb.copy(); nb3[3, 2] = -1  # a padded search label: long -> size_t, position out of range, skipped
    with pytest.raises(drm.DrmError, match="run past"):
        drm.l2_plan_host(1, nb3, stride, k, k_clusters, bound)
    # static dense with an out-of-range id: the reference's mapping keeps original positions and trips its own check
    nb4 = nb.copy(); nb4[2, 1] = bound + 5
    with pytest.raises(drm.DrmError, match="Invalid mapping index"):
        drm.l2_plan_host(0, nb4, 1, 5, 5, bound)


def test_direct_ids_plan_in_parallel_equals_the_indexed_plan():
    """More than 65 536 ids: the direct-ids expansion runs its fill loop on all host threads; skipped ids (position beyond the
    bound) and clipped ones (first windows of the genome) shift the flat list exactly as in the indexed plan."""
    import deepreadmapper_b200 as drm
    g = np.random.default_rng(17)
    n, k_in, stride, kc, k, bound = 3000, 32, 8, 32, 64, 60000
    nb = g.integers(1, bound // stride + 40, (n, k_in)).astype(np.int64)  # some ids expand beyond the bound: skipped
    nb[::7, 3] = 0                                                          # clipped at the start
    try:
        a = drm.l2_plan_host(1, nb, stride, k, kc, bound)
    except drm.DrmError as e:  # both forms refuse the same inputs
        with pytest.raises(drm.DrmError):
            drm.l2_plan_host(1, nb, stride, k, kc, bound, direct_ids=True)
        assert "run past" in str(e)
        nb = g.integers(2, bound // stride - 2, (n, k_in)).astype(np.int64)
        a = drm.l2_plan_host(1, nb, stride, k, kc, bound)
    b = drm.l2_plan_host(1, nb, stride, k, kc, bound, direct_ids=True)
    assert len(b["unique"]) == 0 and np.array_equal(b["map"], a["unique"][a["map"]].astype(np.uint32)) and np.array_equal(b["starts"], a["starts"])
