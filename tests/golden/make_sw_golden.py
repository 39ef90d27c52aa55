"""Regenerates tests/golden/sw_golden.json by running the UNMODIFIED reference (oracle/_ref, compiled
from /root/reference in place).  Run in the build container only:  python tests/golden/make_sw_golden.py

Contents
  fixtures    the reference's own small test inputs, carried here because /root/reference does not
              exist on the GPU box: genome of tests/ecoli_150.fna (as extract_FASTA_sequence returns
              it), the read sequences / ids of tests/test_data.fastq (as preprocess_fastq returns
              them, '<'..'>' decorated) and the lines of tests/test_data_quer.txt.
  sw_pairs100 the src/test_sw_score.cpp batch workload: calc_sw_score(seq[i], seq[i+1]), i<100.
  sw_named    single known answers quoted in SURVEY.md §4.
  sw_matrix   calc_sw_score(window w, read r) for every window of the 1702-window table x 16 reads
              (row sums + full matrix for the first 4 reads).
  rerank      sw_reranker / post_process_sw_{static,dynamic} outputs for dense and sparse settings.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import bindings as ob  # noqa: E402

REF = "/root/reference/tests"


def main():
    import ctypes as C
    R = ob.ref()
    # ---- fixtures through the reference's own parsers
    cap = 1 << 20
    buf = C.create_string_buffer(cap)
    glen = R.ref_extract_fasta_sequence(f"{REF}/ecoli_150.fna".encode(), buf, cap)
    genome = buf.raw[:glen].decode()

    n = C.c_int64(); sb = C.c_int64(); ib = C.c_int64()
    R.ref_preprocess_fastq(f"{REF}/test_data.fastq".encode(), None, None, None, None, C.byref(n), C.byref(sb), C.byref(ib))
    sflat = C.create_string_buffer(sb.value + 1); iflat = C.create_string_buffer(ib.value + 1)
    soff = np.zeros(n.value + 1, dtype=np.int64); ioff = np.zeros(n.value + 1, dtype=np.int64)
    R.ref_preprocess_fastq(f"{REF}/test_data.fastq".encode(), sflat, soff.ctypes.data, iflat, ioff.ctypes.data,
                           C.byref(n), C.byref(sb), C.byref(ib))
    reads = [sflat.raw[soff[i]:soff[i + 1]].decode() for i in range(n.value)]
    read_ids = [iflat.raw[ioff[i]:ioff[i + 1]].decode() for i in range(n.value)]
    quer = [l for l in open(f"{REF}/test_data_quer.txt").read().split("\n") if l]

    # window table through the reference's FASTA path (lookup mode) — must equal test_data_ref.txt
    nw = C.c_int64(); wb = C.c_int64()
    R.ref_preprocess_fasta(f"{REF}/ecoli_150.fna".encode(), 150, 1, 1, None, None, None, C.byref(nw), C.byref(wb))
    wflat = C.create_string_buffer(wb.value + 1); woff = np.zeros(nw.value + 1, dtype=np.int64)
    wlab = np.zeros(nw.value, dtype=np.uint64)
    R.ref_preprocess_fasta(f"{REF}/ecoli_150.fna".encode(), 150, 1, 1, wflat, woff.ctypes.data, wlab.ctypes.data,
                           C.byref(nw), C.byref(wb))
    windows = [wflat.raw[woff[i]:woff[i + 1]].decode() for i in range(nw.value)]
    ref_txt = [l for l in open(f"{REF}/test_data_ref.txt").read().split("\n") if l]
    assert windows == ref_txt, "format_fasta(lookup) != tests/test_data_ref.txt"
    assert list(wlab) == list(range(nw.value))

    def sw(a, b):
        a = a.encode(); b = b.encode()
        return R.ref_calc_sw_score(a, len(a), b, len(b))

    g = {"fixtures": {"genome": genome, "reads": reads, "read_ids": read_ids, "quer": quer,
                      "n_windows": nw.value, "ref_len": 150}}
    g["sw_pairs100"] = [sw(quer[i], quer[i + 1]) for i in range(100)]
    q0 = quer[0]
    g["sw_named"] = {
        "q0_q0": sw(q0, q0), "w561_q0": sw(windows[561], q0), "w560_q0": sw(windows[560], q0),
        "w561_q0_decorated": sw(windows[561], "<" + q0 + ">"),
        "empty_q0": sw("", q0), "q0_empty": sw(q0, ""), "N_vs_N": sw("NNNN", "NNNN"), "lower_vs_upper": sw("acgt", "ACGT"),
        "ragged": sw(windows[10][:97], q0[13:140]),
    }
    nq = 16
    mat = np.zeros((nq, nw.value), dtype=np.int32)
    for r in range(nq):
        mat[r] = ob.sw_pairs(windows, [quer[r]] * nw.value, lib=R)
    g["sw_matrix"] = {"n_reads": nq, "row_sums": mat.sum(1).tolist(), "argmax": mat.argmax(1).tolist(),
                      "max": mat.max(1).tolist(), "rows4": mat[:4].tolist(), "checksum": int(mat.sum())}
    assert g["sw_matrix"]["checksum"] == 786674, g["sw_matrix"]["checksum"]

    # ---- rerank goldens: neighbours are synthetic but deterministic
    rng = np.random.default_rng(20261104)
    win_flat = "".join(windows).encode()
    cases = []

    def run(mode, nb, stride, k, kc, queries, tag):
        refb, nref = (win_flat, nw.value) if mode == 0 else (genome.encode(), len(genome))
        case = {"tag": tag, "mode": mode, "stride": stride, "k": k, "k_clusters": kc, "neighbors": nb.tolist(),
                "queries": queries}
        # sw_reranker throws from inside the reference's `omp parallel for` (post_processor.cpp:394,491), which
        # std::terminate()s the process; so queries that would throw are fed to sw_reranker directly instead.
        take = min(kc, nb.shape[1])
        bound = nref
        cand = [ob.expand_ids(mode, nb[i, :take].astype(np.uint64), stride, bound) for i in range(len(queries))]
        short = [i for i, c in enumerate(cand) if len(c) < k]
        if short and k <= kc * 2 * stride:
            i = short[0]
            seqs = []
            for cid in cand[i]:
                if mode == 0:
                    seqs.append(windows[int(cid)])
                else:
                    out = C.create_string_buffer(151)
                    ln = R.ref_find_sequence(genome.encode(), len(genome), int(cid), 150, out)
                    seqs.append(out.raw[:ln].decode())
            flat, off = ob.flatten(seqs)
            sc = np.zeros(max(k, 1), dtype=np.int32); ids = np.zeros(max(k, 1), dtype=np.uint64)
            q = queries[i].encode()
            rc = R.ref_sw_reranker(flat, off, len(seqs), np.ascontiguousarray(cand[i], dtype=np.uint64), q, len(q), k, sc, ids)
            assert rc == -1
            case["error"] = R.ref_last_error().decode(); case["error_query"] = i
            case["n_cands"] = [len(c) for c in cand]
        else:
            rc, s, ids, _ = ob.post_process_sw(mode, nb, refb, nref, 150, queries, stride, k, kc, use_ref=True)
            if rc != 0:
                case["error"] = s
            else:
                case["scores"] = s.tolist(); case["ids"] = [[int(v) for v in row] for row in ids]
                case["n_cands"] = [len(c) for c in cand]
        cases.append(case)

    qs = reads[:6]
    # dense, K = 128: every query gets 128 distinct window ids, truth included
    nb = np.stack([rng.permutation(nw.value)[:128] for _ in qs]).astype(np.int64)
    run(0, nb, 1, 128, 128, qs, "dense_static_k128")
    run(1, nb, 1, 128, 128, qs, "dense_dynamic_k128")
    # dense with -1 padding and out-of-range ids: static drops them (-> not enough candidates), dynamic keeps them
    nb2 = nb.copy(); nb2[:, -3:] = -1; nb2[0, 5] = nw.value + 7
    run(0, nb2, 1, 128, 128, qs, "dense_static_padded_error")
    run(1, nb2, 1, 128, 128, qs, "dense_dynamic_padded")
    # NOT run through the reference: static stride==1 with a dropped id AND enough candidates left indexes
    # unique_seqs[] with positions of the uncompacted list (post_processor.cpp:228-233 vs :504-508) -> out-of-bounds
    # read (it segfaults here).  The oracle / CUDA path define the evident intent (drop the id); see DESIGN.md.
    # dense, k < candidates, heavy score ties (tie order of std::partial_sort)
    run(0, nb, 1, 17, 128, qs, "dense_static_k17")
    run(0, nb[:, :40], 1, 8, 40, qs, "dense_static_kc40_k8")
    # sparse: stride 8 ids expanded in id space; static bound = #windows, dynamic bound = genome length
    nsp = nw.value // 8
    nbs = np.stack([rng.permutation(nsp + 4)[:32] for _ in qs]).astype(np.int64)  # a few beyond the bound
    run(0, nbs, 8, 128, 32, qs, "sparse8_static_k128")
    run(1, nbs, 8, 128, 32, qs, "sparse8_dynamic_k128")
    run(0, nbs, 8, 300, 32, qs, "sparse8_static_k300")
    run(0, nbs[:, :5], 8, 20, 5, qs, "sparse8_static_default_kclusters")
    run(0, nbs, 8, 600, 32, qs, "sparse8_k_too_large")
    nbs3 = np.stack([rng.permutation(nw.value // 3)[:16] for _ in qs]).astype(np.int64)
    run(0, nbs3, 3, 40, 16, qs, "sparse3_static")
    run(1, nbs3, 3, 40, 16, qs, "sparse3_dynamic")
    # ragged / undecorated / lower-case queries
    odd = [quer[0][:80], quer[1], quer[2].lower(), "<" + quer[3][5:120] + ">", "N" * 150, quer[5][:1]]
    run(0, nb, 1, 64, 128, odd, "dense_static_odd_queries")
    g["rerank"] = cases

    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "sw_golden.json")
    with open(out, "w") as f:
        json.dump(g, f, separators=(",", ":"))
    print("wrote", out, os.path.getsize(out), "bytes;", len(cases), "rerank cases")


if __name__ == "__main__":
    main()
