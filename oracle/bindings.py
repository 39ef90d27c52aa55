"""ctypes bindings for the CPU oracle (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this module.  The product package (deepreadmapper_b200) never does.

  liboracle.so        oracle/sw_oracle.c + oracle/hnswpq_oracle.cpp  (this repo's restatement)
  _ref/libdrm_ref.so  the unmodified reference sources compiled in place + oracle/ref_shim.cpp
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_HERE, "liboracle.so")
REF_SO = os.path.join(_HERE, "_ref", "libdrm_ref.so")

_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")


def build(ref: bool = True) -> None:
    """Compile the oracle (and oracle/_ref when /root/reference is present)."""
    subprocess.check_call(["make", "-s", "-f", os.path.join(_HERE, "Makefile"), ORACLE_SO])
    if ref:
        subprocess.check_call(["make", "-s", "-f", os.path.join(_HERE, "Makefile"), "ref"])


def flatten(seqs):
    """list[bytes|str] -> (flat bytes buffer, int64 offsets[n+1])."""
    bs = [s.encode() if isinstance(s, str) else bytes(s) for s in seqs]
    off = np.zeros(len(bs) + 1, dtype=np.int64)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs])
    return b"".join(bs), off


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        L = C.CDLL(ORACLE_SO)
        L.orc_sw_score.restype = C.c_int
        L.orc_sw_score.argtypes = [C.c_char_p, C.c_int64, C.c_char_p, C.c_int64]
        L.orc_sw_pairs.restype = None
        L.orc_sw_pairs.argtypes = [C.c_char_p, _i64p, C.c_char_p, _i64p, C.c_int64, _i32p, C.c_int]
        L.orc_partial_sort_desc.restype = None
        L.orc_partial_sort_desc.argtypes = [_i32p, C.c_int64, C.c_int64, _i64p]
        L.orc_find_sequence.restype = C.c_int64
        L.orc_find_sequence.argtypes = [C.c_char_p, C.c_int64, C.c_uint64, C.c_int64, C.c_char_p]
        L.orc_expand_ids.restype = C.c_int64
        L.orc_expand_ids.argtypes = [C.c_int, _u64p, C.c_int64, C.c_uint64, C.c_uint64, _u64p]
        L.orc_post_process_sw.restype = C.c_int64
        L.orc_post_process_sw.argtypes = [C.c_int, _i64p, C.c_int64, C.c_int64, C.c_char_p, C.c_int64, C.c_int64,
                                          C.c_char_p, _i64p, C.c_int64, C.c_int64, C.c_int64, _i32p, _u64p, _i64p, C.c_int]
        L.orc_sw_align.restype = None
        L.orc_sw_align.argtypes = [C.c_char_p, C.c_int64, C.c_char_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_last_error.restype = C.c_char_p
        L.orc_index_read.restype = C.c_void_p
        L.orc_index_read.argtypes = [C.c_char_p]
        L.orc_index_free.restype = None
        L.orc_index_free.argtypes = [C.c_void_p]
        L.orc_index_info.restype = None
        L.orc_index_info.argtypes = [C.c_void_p, _i64p]
        L.orc_index_validate.restype = C.c_int
        L.orc_index_validate.argtypes = [C.c_void_p]
        L.orc_index_arrays.restype = None
        L.orc_index_arrays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_search.restype = C.c_int
        L.orc_search.argtypes = [C.c_void_p, _f32p, C.c_int64, C.c_int, C.c_int, C.c_int, _f32p, _i64p, _i64p,
                                 C.c_void_p, C.c_int]
        L.orc_adc_scan.restype = C.c_int
        L.orc_adc_scan.argtypes = [C.c_void_p, _f32p, C.c_int64, C.c_int, _f32p, _i64p, C.c_int]
        L.orc_adc_distances.restype = None
        L.orc_adc_distances.argtypes = [C.c_void_p, _f32p, _i64p, C.c_int64, _f32p]
        _oracle = L
    return _oracle


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def ref():
    global _ref
    if _ref is None:
        L = C.CDLL(REF_SO)
        L.ref_last_error.restype = C.c_char_p
        L.ref_calc_sw_score.restype = C.c_int
        L.ref_calc_sw_score.argtypes = [C.c_char_p, C.c_int64, C.c_char_p, C.c_int64]
        L.ref_sw_pairs.restype = None
        L.ref_sw_pairs.argtypes = [C.c_char_p, _i64p, C.c_char_p, _i64p, C.c_int64, _i32p, C.c_int]
        L.ref_sw_reranker.restype = C.c_int
        L.ref_sw_reranker.argtypes = [C.c_char_p, _i64p, C.c_int64, _u64p, C.c_char_p, C.c_int64, C.c_int64, _i32p, _u64p]
        L.ref_post_process_sw.restype = C.c_int
        L.ref_post_process_sw.argtypes = [C.c_int, _i64p, C.c_int64, C.c_int64, C.c_char_p, C.c_int64, C.c_int64,
                                          C.c_char_p, _i64p, C.c_int64, C.c_int64, C.c_int64, _i32p, _u64p,
                                          C.POINTER(C.c_int64)]
        L.ref_find_sequence.restype = C.c_int64
        L.ref_find_sequence.argtypes = [C.c_char_p, C.c_int64, C.c_uint64, C.c_int64, C.c_char_p]
        L.ref_reverse_complement.restype = None
        L.ref_reverse_complement.argtypes = [C.c_char_p, C.c_int64, C.c_char_p]
        L.ref_preprocess_fasta.restype = C.c_int
        L.ref_preprocess_fasta.argtypes = [C.c_char_p, C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.ref_extract_fasta_sequence.restype = C.c_int64
        L.ref_extract_fasta_sequence.argtypes = [C.c_char_p, C.c_void_p, C.c_int64]
        L.ref_preprocess_fastq.restype = C.c_int
        L.ref_preprocess_fastq.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.ref_save_results.restype = C.c_int
        L.ref_save_results.argtypes = [_i64p, _f32p, C.c_int64, C.c_int64, C.c_int64, C.c_char_p, C.c_char_p]
        if hasattr(L, "ref_write_sam"):
            L.ref_write_sam.restype = C.c_int
            L.ref_write_sam.argtypes = [C.c_char_p, C.c_void_p, C.c_int64, C.c_int64, C.c_char_p, C.c_int, C.c_int64, C.c_char_p]
        L.ref_save_config.restype = C.c_int
        L.ref_save_config.argtypes = [C.c_char_p] + [C.c_uint64] * 8 + [C.c_char_p]
        L.ref_load_config_size.restype = C.c_int64
        L.ref_load_config_size.argtypes = [C.c_char_p, C.c_char_p]
        _ref = L
    return _ref


# ------------------------------------------------------------------ convenience wrappers (oracle)
def sw_score(a, b) -> int:
    a = a.encode() if isinstance(a, str) else a
    b = b.encode() if isinstance(b, str) else b
    return oracle().orc_sw_score(a, len(a), b, len(b))


def sw_pairs(a_list, b_list, threads: int = 0, lib=None):
    a_flat, a_off = flatten(a_list)
    b_flat, b_off = flatten(b_list)
    out = np.zeros(len(a_list), dtype=np.int32)
    threads = threads or (os.cpu_count() or 1)
    if lib is None:
        oracle().orc_sw_pairs(a_flat, a_off, b_flat, b_off, len(a_list), out, threads)
    else:
        lib.ref_sw_pairs(a_flat, a_off, b_flat, b_off, len(a_list), out, threads)
    return out


def partial_sort_desc(scores, k):
    scores = np.ascontiguousarray(scores, dtype=np.int32)
    idx = np.zeros(len(scores), dtype=np.int64)
    oracle().orc_partial_sort_desc(scores, len(scores), k, idx)
    return idx[:k]


def expand_ids(mode, ids, stride, bound):
    ids = np.ascontiguousarray(ids, dtype=np.uint64)
    per = 1 if stride == 1 else 2 * stride - 1
    out = np.zeros(len(ids) * per + 1, dtype=np.uint64)
    m = oracle().orc_expand_ids(mode, ids, len(ids), stride, bound, out)
    return out[:m]


def sw_align(window: bytes, read: bytes, max_ops: int = 64):
    """SW with traceback (the definition in sw_oracle.c: orc_sw_align).  Returns dict(score, ref_beg, ref_end, q_beg, q_end,
    n_ops, nm, ops=[(length, 'M'|'I'|'D')...])."""
    out = np.zeros(7, dtype=np.int32)
    ops = np.zeros(max_ops, dtype=np.uint32)
    oracle().orc_sw_align(window, len(window), read, len(read), out.ctypes.data, ops.ctypes.data, max_ops)
    keys = ("score", "ref_beg", "ref_end", "q_beg", "q_end", "n_ops", "nm")
    d = {k: int(v) for k, v in zip(keys, out)}
    d["ops"] = [(int(w) >> 4, "MID"[int(w) & 15]) for w in ops[:min(d["n_ops"], max_ops)]]
    return d


def post_process_sw(mode, neighbors, ref_bytes, n_ref, ref_len, queries, stride, k, k_clusters, threads=0, use_ref=False):
    """Returns (rc, scores[n,k], ids[n,k], n_cands[n] or None).  rc==0 ok."""
    neighbors = np.ascontiguousarray(neighbors, dtype=np.int64)
    n, k_in = neighbors.shape
    q_flat, q_off = flatten(queries)
    scores = np.zeros((n, k), dtype=np.int32)
    ids = np.zeros((n, k), dtype=np.uint64)
    if use_ref:
        cnt = C.c_int64(0)
        rc = ref().ref_post_process_sw(mode, neighbors, n, k_in, ref_bytes, n_ref, ref_len, q_flat, q_off, stride, k,
                                       k_clusters, scores.reshape(-1), ids.reshape(-1), C.byref(cnt))
        if rc != 0:
            return rc, ref().ref_last_error().decode(), None, None
        assert cnt.value == n * k, (cnt.value, n, k)
        return 0, scores, ids, None
    ncand = np.zeros(n, dtype=np.int64)
    rc = oracle().orc_post_process_sw(mode, neighbors, n, k_in, ref_bytes, n_ref, ref_len, q_flat, q_off, stride, k,
                                      k_clusters, scores.reshape(-1), ids.reshape(-1), ncand,
                                      threads or (os.cpu_count() or 1))
    return rc, scores, ids, ncand


class OracleIndex:
    """Handle on an IHNp file read by the oracle's own reader."""

    def __init__(self, path: str):
        self._h = oracle().orc_index_read(path.encode())
        if not self._h:
            raise RuntimeError(oracle().orc_last_error().decode())
        info = np.zeros(10, dtype=np.int64)
        oracle().orc_index_info(self._h, info)
        (self.d, self.ntotal, self.M_pq, self.nbits, self.M_hnsw, self.entry_point, self.max_level,
         self.efConstruction, self.efSearch, self.n_links) = (int(v) for v in info)

    def close(self):
        if self._h:
            oracle().orc_index_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def validate(self) -> int:
        return oracle().orc_index_validate(self._h)

    def arrays(self):
        """(centroids [M][256][dsub], codes ntotal x M, {degree statistics}) - used by the builder tests."""
        dsub = self.d // self.M_pq
        cent = np.zeros((self.M_pq, 256, dsub), dtype=np.float32)
        codes = np.zeros((self.ntotal, self.M_pq), dtype=np.uint8)
        deg = np.zeros(3, dtype=np.float64)
        oracle().orc_index_arrays(self._h, cent.ctypes.data, codes.ctypes.data, deg.ctypes.data)
        return cent, codes, dict(mean_degree0=float(deg[0]), empty_rows0=float(deg[1]), mean_degree1=float(deg[2]))

    def search(self, x, k, ef, tie_rule=1, threads=0, per_query=False):
        x = np.ascontiguousarray(x, dtype=np.float32)
        n = x.shape[0]
        D = np.zeros((n, k), dtype=np.float32)
        I = np.zeros((n, k), dtype=np.int64)
        stats = np.zeros(4, dtype=np.int64)
        pq = np.zeros((n, 2), dtype=np.int32) if per_query else None
        rc = oracle().orc_search(self._h, x.reshape(-1), n, k, ef, tie_rule, D.reshape(-1), I.reshape(-1), stats,
                                 pq.ctypes.data if pq is not None else None, threads)
        if rc != 0:
            raise RuntimeError(oracle().orc_last_error().decode())
        st = dict(ndis=int(stats[0]), nhops=int(stats[1]), ndis_up=int(stats[2]), nhops_up=int(stats[3]))
        return (D, I, st, pq) if per_query else (D, I, st)

    def adc_scan(self, x, k, threads=0):
        x = np.ascontiguousarray(x, dtype=np.float32)
        n = x.shape[0]
        D = np.zeros((n, k), dtype=np.float32)
        I = np.zeros((n, k), dtype=np.int64)
        oracle().orc_adc_scan(self._h, x.reshape(-1), n, k, D.reshape(-1), I.reshape(-1), threads)
        return D, I

    def adc_distances(self, xq, ids):
        xq = np.ascontiguousarray(xq, dtype=np.float32)
        ids = np.ascontiguousarray(ids, dtype=np.int64)
        out = np.zeros(len(ids), dtype=np.float32)
        oracle().orc_adc_distances(self._h, xq, ids, len(ids), out)
        return out
