"""Pins oracle/sw_oracle.c (the CPU restatement) against the golden vectors generated from the
unmodified reference (tests/golden/sw_golden.json) and, when oracle/_ref is present, against the
compiled reference itself on randomised inputs.  CPU only."""
import ctypes as C

import numpy as np
import pytest


def test_sw_pairs100_matches_reference_test_sw_score(golden, ob):
    # src/test_sw_score.cpp:44-52 batch workload
    q = golden["fixtures"]["quer"]
    got = [ob.sw_score(q[i], q[i + 1]) for i in range(100)]
    assert got == golden["sw_pairs100"]
    assert sum(got) == 2696


def test_sw_named_known_answers(golden, fixtures, ob):
    q0 = fixtures["quer"][0]
    w = fixtures["windows"]
    n = golden["sw_named"]
    assert ob.sw_score(q0, q0) == n["q0_q0"] == 150
    assert ob.sw_score(w[561], q0) == n["w561_q0"] == 138
    assert ob.sw_score(w[560], q0) == n["w560_q0"]
    assert ob.sw_score(w[561], "<" + q0 + ">") == n["w561_q0_decorated"] == 138
    assert ob.sw_score("", q0) == n["empty_q0"] == 0
    assert ob.sw_score(q0, "") == n["q0_empty"] == 0
    assert ob.sw_score("NNNN", "NNNN") == n["N_vs_N"] == 4
    assert ob.sw_score("acgt", "ACGT") == n["lower_vs_upper"] == 0
    assert ob.sw_score(w[10][:97], q0[13:140]) == n["ragged"]


def test_sw_matrix_all_windows(golden, fixtures, ob):
    m = golden["sw_matrix"]
    w = fixtures["windows"]
    tot = 0
    for r in range(m["n_reads"]):
        row = ob.sw_pairs(w, [fixtures["quer"][r]] * len(w))
        assert int(row.sum()) == m["row_sums"][r]
        assert int(row.argmax()) == m["argmax"][r] and int(row.max()) == m["max"][r]
        if r < 4:
            assert row.tolist() == m["rows4"][r]
        tot += int(row.sum())
    assert tot == m["checksum"] == 786674


def _run_case(ob, fixtures, case, **kw):
    if case["mode"] == 0:
        refb, nref = "".join(fixtures["windows"]).encode(), fixtures["n_windows"]
    else:
        refb, nref = fixtures["genome"].encode(), len(fixtures["genome"])
    nb = np.array(case["neighbors"], dtype=np.int64)
    return ob.post_process_sw(case["mode"], nb, refb, nref, 150, case["queries"], case["stride"], case["k"],
                              case["k_clusters"], **kw)


def test_rerank_goldens(golden, fixtures, ob):
    assert len(golden["rerank"]) >= 14
    for case in golden["rerank"]:
        rc, s, ids, nc = _run_case(ob, fixtures, case)
        if "error" in case:
            assert rc < 0, case["tag"]
            if case["error"].startswith("Final k too large"):
                assert rc == -1000000000
            else:
                i = case["error_query"]
                assert f"({nc[i]} < {case['k']})" in case["error"]
            continue
        assert rc == 0, case["tag"]
        assert s.tolist() == case["scores"], case["tag"]
        assert [[int(v) for v in r] for r in ids] == case["ids"], case["tag"]
        assert nc.tolist() == case["n_cands"]


needs_ref = pytest.mark.skipif(not __import__("oracle.bindings", fromlist=["x"]).have_ref(),
                               reason="oracle/_ref not built (no /root/reference)")


@needs_ref
def test_oracle_vs_compiled_reference_random_pairs(ob):
    rng = np.random.default_rng(5)
    alpha = np.frombuffer(b"ACGTN", dtype=np.uint8)
    a, b = [], []
    for _ in range(400):
        la, lb = rng.integers(0, 200, 2)
        x = alpha[rng.integers(0, 5 if rng.random() < .3 else 4, la)].tobytes()
        if rng.random() < .5 and la > 10:   # make b a mutated copy so that scores are non-trivial
            y = bytearray(x[: max(1, lb)])
            for p in rng.integers(0, len(y), max(1, len(y) // 12)):
                y[p] = alpha[rng.integers(0, 4)]
            y = bytes(y)
        else:
            y = alpha[rng.integers(0, 4, lb)].tobytes()
        if rng.random() < .1:
            y = y.lower()
        a.append(x); b.append(y)
    mine = ob.sw_pairs(a, b, threads=2)
    theirs = ob.sw_pairs(a, b, threads=2, lib=ob.ref())
    assert mine.tolist() == theirs.tolist()


@needs_ref
def test_partial_sort_tie_order_vs_libstdcxx(ob):
    """sw_reranker's std::partial_sort is not stable; the restatement must reproduce its tie order."""
    R = ob.ref()
    rng = np.random.default_rng(11)
    for trial in range(300):
        n = int(rng.integers(1, 520))
        k = int(rng.integers(1, n + 1))
        hi = int(rng.choice([2, 5, 30, 150]))
        # candidate i is a run of i%hi matching bases: score == i%hi  => controlled, tie-heavy scores
        q = "A" * 200
        want_scores = rng.integers(0, hi, n)
        cands = ["A" * int(s) + "C" for s in want_scores]
        flat, off = ob.flatten(cands)
        ids = np.arange(1000, 1000 + n, dtype=np.uint64)
        sc = np.zeros(k, dtype=np.int32); out_ids = np.zeros(k, dtype=np.uint64)
        rc = R.ref_sw_reranker(flat, off, n, ids, q.encode(), len(q), k, sc, out_ids)
        assert rc == k
        idx = ob.partial_sort_desc(want_scores.astype(np.int32), k)
        assert (ids[idx]).tolist() == out_ids.tolist(), (trial, n, k, hi)
        assert want_scores[idx].tolist() == sc.tolist()


@needs_ref
def test_expansion_and_window_fetch_vs_reference(fixtures, ob):
    R = ob.ref()
    g = fixtures["genome"].encode()
    out_r = C.create_string_buffer(400); out_o = C.create_string_buffer(400)
    for idv in [0, 1, 2, 561, 1700, 1701, 1702, 1703, 1999, 2**64 - 1, 2**63]:
        for L in (150, 31, 1):
            lr = R.ref_find_sequence(g, len(g), idv, L, out_r)
            lo = ob.oracle().orc_find_sequence(g, len(g), idv, L, out_o)
            assert lr == lo and out_r.raw[:lr] == out_o.raw[:lo], (idv, L)
    # whole-stage agreement on random neighbour tables (no short/invalid queries: those terminate the reference)
    rng = np.random.default_rng(3)
    reads = fixtures["reads"][:5]
    for mode in (0, 1):
        for stride, kc, k in [(1, 64, 64), (1, 64, 10), (4, 20, 50), (8, 32, 128), (8, 5, 20), (16, 8, 60)]:
            hi = fixtures["n_windows"] // stride if mode == 0 else len(fixtures["genome"]) // stride
            hi = max(hi - (1 if mode == 1 else 0), kc + 1)
            nb = np.stack([rng.permutation(hi)[:kc] for _ in reads]).astype(np.int64)
            refb, nref = ("".join(fixtures["windows"]).encode(), fixtures["n_windows"]) if mode == 0 else (g, len(g))
            rc1, s1, i1, nc = ob.post_process_sw(mode, nb, refb, nref, 150, reads, stride, k, kc)
            if rc1 != 0:
                continue
            rc2, s2, i2, _ = ob.post_process_sw(mode, nb, refb, nref, 150, reads, stride, k, kc, use_ref=True)
            assert rc2 == 0
            assert s1.tolist() == s2.tolist() and i1.tolist() == i2.tolist(), (mode, stride, kc, k)


def test_traceback_oracle_is_pinned_by_the_score(ob, fixtures):
    """orc_sw_align is a definition (the reference has no traceback): its score must equal calc_sw_score and re-scoring the
    reported operations along the reported intervals must give that score back."""
    rng = np.random.default_rng(4)
    pairs = [(fixtures["windows"][i].encode(), fixtures["reads"][i % len(fixtures["reads"])].strip("<>").encode()) for i in range(0, 400, 7)]
    for t in range(200):
        la, lb = int(rng.integers(0, 80)), int(rng.integers(0, 80))
        a = bytes(rng.choice(list(b"ACGTN"), la))
        b = bytearray(a[:lb]) if t % 2 else bytearray(rng.choice(list(b"ACGT"), lb))
        for _ in range(len(b) // 7):
            b[int(rng.integers(0, len(b)))] = ord("C")
        if t % 5 == 0 and len(b) > 10:
            del b[5]
        pairs.append((a, bytes(b)))
    for a, b in pairs:
        r = ob.sw_align(a, b, 256)
        assert r["score"] == ob.sw_score(a, b)
        if ob.have_ref():
            assert r["score"] == ob.ref().ref_calc_sw_score(a, len(a), b, len(b))
        i, j, s, nm = r["ref_beg"], r["q_beg"], 0, 0
        for length, op in r["ops"]:
            for _ in range(length):
                if op == "M":
                    s += 1 if a[i] == b[j] else -1
                    nm += a[i] != b[j]
                    i += 1
                    j += 1
                elif op == "D":
                    s, nm, i = s - 1, nm + 1, i + 1
                else:
                    s, nm, j = s - 1, nm + 1, j + 1
        assert (i, j, s, nm) == (r["ref_end"], r["q_end"], r["score"], r["nm"]), (a, b, r)
