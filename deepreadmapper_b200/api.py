"""ctypes mirror of the reference's hot-path call sites over libdrm_cuda.so (include/drm_b200.h).

Names and argument meaning follow the reference (paths relative to the reference repository):
  faiss_search(index, query_data, k, ef)            src/hnswpq/search.cpp:6-57
  post_process_sw_static / post_process_sw_dynamic  src/utils/post_processor.cpp:357-549
  sw_reranker(cand_seqs, cand_ids, query_seq, k)    src/utils/reranker.cpp:3-51
  calc_sw_score(seq1, seq2)                         src/utils/metrics.cpp:10-45
Errors the reference throws as std::runtime_error are raised as DrmError with the same message.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "lib", "libdrm_cuda.so")
_lib = None


class DrmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


class SearchStats(C.Structure):
    _fields_ = [("nhops0", C.c_int64), ("nhops_up", C.c_int64), ("ndis_eval", C.c_int64),
                ("device_ms", C.c_double), ("h2d_ms", C.c_double), ("d2h_ms", C.c_double)]


EXPORTS = ["drm_last_error", "drm_device_count", "drm_version", "drm_index_open", "drm_index_close", "drm_index_info",
           "drm_search", "drm_search_device", "drm_ref_open_genome", "drm_ref_open_windows", "drm_ref_close",
           "drm_sw_pairs", "drm_sw_rerank", "drm_sw_rerank_device", "drm_merge_topk_device", "drm_build_index",
           "drm_search_rerank", "drm_build_index_gpu", "drm_build_index_gpu_device", "drm_fasta_clean", "drm_ref_open_fasta",
           "drm_reads_open_fastq", "drm_reads_close", "drm_reads_info", "drm_reads_copy_out", "drm_reads_device",
           "drm_search_rerank_reads", "drm_sw_align", "drm_embedder_open_ir", "drm_embedder_open_raw", "drm_embedder_close",
           "drm_embedder_ir_info", "drm_tokenize", "drm_embed", "drm_embed_reads", "drm_embed_windows", "drm_embed_device",
           "drm_embed_windows_device", "drm_l2_rerank", "drm_l2_plan_host", "drm_l2_select_host",
           "drm_shards_open", "drm_shards_unique_id", "drm_shards_open_rank", "drm_shards_close", "drm_shards_info", "drm_shards_search_rerank",
           "drm_shards_step_device", "drm_shards_stats", "drm_l2_last_timing", "drm_sw_pairs_device_noalloc"]


class ShardStats(C.Structure):
    _fields_ = [("search_ms", C.c_double), ("exchange_ms", C.c_double), ("merge_ms", C.c_double), ("rerank_ms", C.c_double),
                ("nhops0", C.c_int64), ("nhops_up", C.c_int64), ("ndis_eval", C.c_int64), ("bytes_received", C.c_int64),
                ("steps", C.c_int64)]


XCHG_ALLGATHER, XCHG_ALLTOALL, XCHG_FUSED = 0, 1, 2


def lib_path() -> str:
    return _LIB_PATH


def lib():
    """Load libdrm_cuda.so.  Fails loudly when it has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise DrmError(-3, f"{_LIB_PATH} is missing: run `make lib` (or __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(_LIB_PATH)
        vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
        L.drm_last_error.restype = C.c_char_p
        L.drm_version.restype = C.c_char_p
        L.drm_device_count.restype = i32
        L.drm_index_open.argtypes = [C.c_char_p, C.POINTER(i32), i32, C.POINTER(vp)]
        L.drm_index_close.argtypes = [vp]
        L.drm_index_close.restype = None
        L.drm_index_info.argtypes = [vp, vp]
        L.drm_search.argtypes = [vp, vp, i64, i32, i32, i32, vp, vp, C.POINTER(SearchStats)]
        L.drm_search_device.argtypes = [vp, i32, vp, i64, i32, i32, i32, vp, vp, vp, vp]
        L.drm_ref_open_genome.argtypes = [C.c_char_p, i64, i32, C.POINTER(i32), i32, C.POINTER(vp)]
        L.drm_ref_open_windows.argtypes = [C.c_char_p, i64, i32, C.POINTER(i32), i32, C.POINTER(vp)]
        L.drm_ref_close.argtypes = [vp]
        L.drm_ref_close.restype = None
        L.drm_sw_pairs.argtypes = [C.c_char_p, vp, C.c_char_p, vp, i64, vp]
        L.drm_sw_pairs_device_noalloc.argtypes = [i32, vp, vp, vp, vp, i64, i32, vp, vp]
        L.drm_sw_rerank.argtypes = [vp, i32, vp, i64, i64, C.c_char_p, vp, i64, i64, i64, vp, vp, vp, C.POINTER(C.c_double)]
        L.drm_sw_rerank_device.argtypes = [vp, i32, i32, vp, i64, i64, vp, vp, i32, i64, i64, i64, vp, vp, vp, vp, vp, vp]
        L.drm_merge_topk_device.argtypes = [i32, vp, vp, i32, i64, i32, vp, vp, vp]
        L.drm_build_index.argtypes = [vp, i64, i32, vp, i64, i32, i32, i32, i32, i32, C.c_uint32, C.c_char_p]
        L.drm_sw_align.argtypes = [vp, vp, i64, i32, vp, vp, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]
        L.drm_fasta_clean.argtypes = [vp, i64, i32, vp, i64, C.POINTER(i64)]
        L.drm_ref_open_fasta.argtypes = [vp, i64, i32, C.POINTER(i32), i32, C.POINTER(vp), C.POINTER(i64)]
        L.drm_reads_open_fastq.argtypes = [vp, i64, C.POINTER(i32), i32, C.POINTER(vp)]
        L.drm_reads_close.argtypes = [vp]
        L.drm_reads_close.restype = None
        L.drm_reads_info.argtypes = [vp, vp]
        L.drm_reads_copy_out.argtypes = [vp, vp, vp]
        L.drm_reads_device.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(vp)]
        L.drm_search_rerank_reads.argtypes = [vp, vp, i32, vp, i64, i32, i32, i32, vp, i64, i64, i64, vp, vp, vp, vp,
                                              C.POINTER(SearchStats), vp]
        L.drm_build_index_gpu.argtypes = [vp, i64, i32, vp, i64, i32, i32, i32, i32, i32, C.c_uint32, C.c_char_p, vp]
        L.drm_build_index_gpu_device.argtypes = [vp, i64, i32, i64, i32, i32, i32, i32, i32, C.c_uint32, C.c_char_p, vp]
        L.drm_search_rerank.argtypes = [vp, vp, i32, vp, i64, i32, i32, i32, vp, vp, i64, i64, i64, vp, vp, vp, vp,
                                        C.POINTER(SearchStats), vp]
        L.drm_embedder_open_ir.argtypes = [C.c_char_p, C.c_char_p, i32, C.POINTER(vp)]
        L.drm_embedder_open_raw.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp, i32, i32, C.POINTER(vp)]
        L.drm_embedder_close.argtypes = [vp]
        L.drm_embedder_close.restype = None
        L.drm_embedder_ir_info.argtypes = [C.c_char_p, C.c_char_p, vp, vp]
        L.drm_tokenize.argtypes = [C.c_char_p, vp, i64, i32, vp]
        L.drm_embed.argtypes = [vp, C.c_char_p, vp, i64, vp]
        L.drm_embed_reads.argtypes = [vp, vp, i64, i64, vp]
        L.drm_embed_windows.argtypes = [vp, vp, C.c_uint64, C.c_uint64, i64, vp]
        L.drm_embed_device.argtypes = [vp, vp, vp, i64, i64, vp, vp]
        L.drm_embed_windows_device.argtypes = [vp, vp, C.c_uint64, C.c_uint64, i64, vp, vp]
        L.drm_l2_rerank.argtypes = [vp, vp, i32, vp, vp, i64, i64, vp, i64, i64, i64, vp, vp, C.POINTER(i64)]
        L.drm_l2_plan_host.argtypes = [i32, vp, i64, i64, i64, i64, i64, C.c_uint64, vp, vp, vp, vp]
        L.drm_l2_select_host.argtypes = [vp, i64, i64, vp]
        L.drm_l2_last_timing.argtypes = [vp]
        L.drm_shards_open.argtypes = [C.POINTER(C.c_char_p), vp, C.POINTER(i32), i32, C.POINTER(vp)]
        L.drm_shards_unique_id.argtypes = [vp]
        L.drm_shards_open_rank.argtypes = [C.c_char_p, i64, i32, i32, i32, vp, C.POINTER(vp)]
        L.drm_shards_close.argtypes = [vp]
        L.drm_shards_close.restype = None
        L.drm_shards_info.argtypes = [vp, vp]
        L.drm_shards_search_rerank.argtypes = [vp, vp, i32, vp, i64, i32, i32, i32, vp, vp, i64, i64, i64, i32, i64, vp, vp, vp, vp,
                                               C.POINTER(ShardStats)]
        L.drm_shards_step_device.argtypes = [vp, i32, vp, i64, i32, i32, i32, i32, vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(i64),
                                             C.POINTER(i64), C.POINTER(vp)]
        L.drm_shards_stats.argtypes = [vp, i32, i32, C.POINTER(ShardStats)]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise DrmError(rc, lib().drm_last_error().decode())


def _devs(devices):
    devices = [0] if devices is None else list(devices)
    return (C.c_int * len(devices))(*devices), len(devices)


def flatten(seqs):
    bs = [s.encode() if isinstance(s, str) else bytes(s) for s in seqs]
    off = np.zeros(len(bs) + 1, dtype=np.int64)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs])
    return b"".join(bs), off


class Index:
    """Handle on an IHNp index resident on one or more GPUs (faiss::read_index + cast, src/main.cpp:236-237)."""

    def __init__(self, index_file: str, devices=None):
        arr, n = _devs(devices)
        h = C.c_void_p()
        _check(lib().drm_index_open(os.fsencode(index_file), arr, n, C.byref(h)))
        self._h = h
        info = np.zeros(12, dtype=np.int64)
        _check(lib().drm_index_info(self._h, info.ctypes.data))
        (self.d, self.ntotal, self.M_pq, self.nbits, self.M_hnsw, self.entry_point, self.max_level,
         self.efConstruction, self.efSearch, self.n_links, self.n_replicas, self.device_bytes) = (int(v) for v in info)

    def close(self):
        if getattr(self, "_h", None):
            lib().drm_index_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def search(self, x, k, ef, return_stats=False):
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 2:
            raise DrmError(-1, "Error: Expected 2D array")
        n, d = x.shape
        D = np.empty((n, k), dtype=np.float32)
        I = np.empty((n, k), dtype=np.int64)
        st = SearchStats()
        _check(lib().drm_search(self._h, x.ctypes.data, n, d, k, ef, D.ctypes.data, I.ctypes.data, C.byref(st)))
        if return_stats:
            return D, I, {f: getattr(st, f) for f, _ in SearchStats._fields_}
        return D, I

    def search_device(self, q_ptr, n, k, ef, D_ptr, I_ptr, counters_ptr=0, stream=0, slot=0):
        _check(lib().drm_search_device(self._h, slot, q_ptr, n, self.d, k, ef, D_ptr, I_ptr, counters_ptr or None,
                                       stream or None))


def faiss_search(index: Index, query_data, k=128, ef=128):
    """(neighbors, distances) exactly like the reference's faiss_search: labels int64 (-1 = unfilled)."""
    q = np.asarray(query_data, dtype=np.float32)
    if q.size == 0:
        raise DrmError(-7, "Query data is empty")
    D, I = index.search(q, k, ef)
    return I, D


class Reference:
    """Reference windows on the GPU: a genome (dynamic fetch, id -> pos/strand) or a window table (static)."""

    def __init__(self, genome=None, windows=None, ref_len=150, devices=None, fasta=None):
        arr, n = _devs(devices)
        h = C.c_void_p()
        self.ref_len = ref_len
        if fasta is not None:  # the FASTA file's bytes: cleaned on the device (extract_FASTA_sequence)
            data = bytes(fasta)
            glen = C.c_int64(0)
            buf = C.create_string_buffer(data, len(data)) if data else C.create_string_buffer(1)
            _check(lib().drm_ref_open_fasta(C.addressof(buf), len(data), ref_len, arr, n, C.byref(h), C.byref(glen)))
            self.form, self.length = 0, glen.value
        elif genome is not None:
            g = genome.encode() if isinstance(genome, str) else bytes(genome)
            self.form, self.length = 0, len(g)
            _check(lib().drm_ref_open_genome(g, len(g), ref_len, arr, n, C.byref(h)))
        else:
            flat = b"".join(w.encode() if isinstance(w, str) else bytes(w) for w in windows) if not isinstance(windows, (bytes, bytearray)) else bytes(windows)
            nw = len(flat) // ref_len
            self.form, self.length = 1, nw
            _check(lib().drm_ref_open_windows(flat, nw, ref_len, arr, n, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().drm_ref_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def rerank(self, mode, neighbors, query_seqs, stride, k, k_clusters, return_extra=False):
        nb = np.ascontiguousarray(neighbors, dtype=np.int64)
        n, k_in = nb.shape
        q_flat, q_off = flatten(query_seqs)
        scores = np.zeros((n, k), dtype=np.int32)
        ids = np.zeros((n, k), dtype=np.uint64)
        ncand = np.zeros(n, dtype=np.int32)
        ms = C.c_double(0)
        _check(lib().drm_sw_rerank(self._h, mode, nb.ctypes.data, n, k_in, q_flat, q_off.ctypes.data, stride, k, k_clusters,
                                   scores.ctypes.data, ids.ctypes.data, ncand.ctypes.data, C.byref(ms)))
        if return_extra:
            return scores, ids, ncand, ms.value
        return scores, ids

    def align(self, ids, query_seqs, strip=0, max_ops=64):
        """SW with traceback for (read i, window ids[i, s]) - the real CIGAR behind a rerank score (drm_sw_align).
        ids: (n, per_read) window ids.  Returns dict of int32 arrays shaped (n, per_read): score, ref_beg, ref_end, q_beg,
        q_end, n_ops, nm, and ops (n, per_read, max_ops) uint32 words (length << 4 | op; 0 = M, 1 = I, 2 = D)."""
        ids = np.ascontiguousarray(ids, dtype=np.uint64)
        if ids.ndim == 1:
            ids = ids.reshape(-1, 1)
        n, per = ids.shape
        q_flat, q_off = flatten(query_seqs)
        out = {k: np.zeros((n, per), dtype=np.int32) for k in ("score", "ref_beg", "ref_end", "q_beg", "q_end", "n_ops", "nm")}
        ops = np.zeros((n, per, max_ops), dtype=np.uint32)
        _check(lib().drm_sw_align(self._h, ids.ctypes.data, n, per, q_flat, q_off.ctypes.data, strip, max_ops,
                                  out["score"].ctypes.data, out["ref_beg"].ctypes.data, out["ref_end"].ctypes.data,
                                  out["q_beg"].ctypes.data, out["q_end"].ctypes.data, out["n_ops"].ctypes.data,
                                  out["nm"].ctypes.data, ops.ctypes.data))
        out["ops"] = ops
        return out

    def rerank_device(self, mode, nb_ptr, n, k_in, qflat_ptr, qoff_ptr, max_qlen, stride, k, k_clusters, scores_ptr, ids_ptr,
                      ncand_ptr=0, min_ptr=0, cells_ptr=0, stream=0, slot=0):
        _check(lib().drm_sw_rerank_device(self._h, slot, mode, nb_ptr, n, k_in, qflat_ptr, qoff_ptr, max_qlen, stride, k,
                                          k_clusters, scores_ptr, ids_ptr, ncand_ptr or None, min_ptr or None,
                                          cells_ptr or None, stream or None))


def _post_process(mode, neighbors, ref: Reference, query_seqs, ref_len, stride, k, k_clusters):
    if ref.ref_len != ref_len:
        raise DrmError(-1, "ref_len does not match the Reference handle")
    scores, ids = ref.rerank(mode, neighbors, query_seqs, stride, k, k_clusters)
    return scores.reshape(-1), ids.reshape(-1)


def post_process_sw_static(neighbors, distances, ref: Reference, query_seqs, ref_len, stride, k, k_clusters=5):
    """(scores[n*k], ids[n*k]) — the int / size_t halves of the reference's returned tuple (sequences are
    recoverable from ids; strings never leave the host)."""
    return _post_process(0, neighbors, ref, query_seqs, ref_len, stride, k, k_clusters)


def post_process_sw_dynamic(neighbors, distances, ref: Reference, query_seqs, ref_len, stride, k, k_clusters=5):
    return _post_process(1, neighbors, ref, query_seqs, ref_len, stride, k, k_clusters)


def sw_pairs(a_list, b_list):
    a_flat, a_off = flatten(a_list)
    b_flat, b_off = flatten(b_list)
    out = np.zeros(len(a_list), dtype=np.int32)
    _check(lib().drm_sw_pairs(a_flat, a_off.ctypes.data, b_flat, b_off.ctypes.data, len(a_list), out.ctypes.data))
    return out


SW_PAIRS_NOALLOC_MAX_LEN_B = 4095
SW_PAIRS_NOALLOC_SCRATCH_BYTES = 16 << 20


def sw_pairs_device_noalloc(device, a_flat_ptr, a_off_ptr, b_flat_ptr, b_off_ptr, n, max_len_b, scores_ptr, stream=0):
    """drm_sw_pairs on device buffers without any allocation inside the call (static 16 MiB scratch per GPU; calls that
    overlap on one GPU race).  No counterpart in the reference."""
    _check(lib().drm_sw_pairs_device_noalloc(device, a_flat_ptr, a_off_ptr, b_flat_ptr, b_off_ptr, n, max_len_b, scores_ptr, stream or None))


def calc_sw_score(seq1, seq2) -> int:
    return int(sw_pairs([seq1], [seq2])[0])


def sw_reranker(cand_seqs, cand_ids, query_seq, k):
    """(top_scores, top_ids) for one query over explicit candidate strings (src/utils/reranker.cpp:3-51)."""
    n = len(cand_seqs)
    if n == 0 or k == 0:
        return np.zeros(0, np.int32), np.zeros(0, np.uint64)
    if n < k:
        raise DrmError(-5, f"Not enough candidates ({n} < {k})")
    lens = {len(c) for c in cand_seqs}
    if len(lens) != 1:
        raise DrmError(-1, "sw_reranker needs equal-length candidate windows")
    ref = Reference(windows=list(cand_seqs), ref_len=lens.pop())
    try:
        nb = np.arange(n, dtype=np.int64)[None, :]
        scores, idx = ref.rerank(0, nb, [query_seq], 1, k, n)
    finally:
        ref.close()
    return scores[0], np.asarray(cand_ids, dtype=np.uint64)[idx[0].astype(np.int64)]


def search_rerank(index: Index, ref: Reference, mode, q, q_flat_ptr, q_off, stride, k, k_clusters, k_search, ef,
                  D=None, I=None, sw_scores=None, sw_ids=None):
    """The pipeline's search + SW-rerank stage in one C-ABI call on HOST buffers (src/main.cpp:270-354).
    q: (n, d) float32 ndarray; q_flat_ptr: address (or bytes) of the concatenated read sequences; q_off: int64
    ndarray (n+1).  Output arrays may be preallocated (e.g. pinned) ndarrays; missing ones are created.
    Returns (D, I, sw_scores, sw_ids, stats dict, [device_span_ms, wall_ms])."""
    n, d = q.shape
    D = np.empty((n, k_search), np.float32) if D is None else D
    I = np.empty((n, k_search), np.int64) if I is None else I
    sw_scores = np.empty((n, k), np.int32) if sw_scores is None else sw_scores
    sw_ids = np.empty((n, k), np.uint64) if sw_ids is None else sw_ids
    st = SearchStats()
    tm = np.zeros(2, dtype=np.float64)
    qf = q_flat_ptr if isinstance(q_flat_ptr, int) else C.cast(C.c_char_p(q_flat_ptr), C.c_void_p).value
    _check(lib().drm_search_rerank(index._h, ref._h, mode, q.ctypes.data, n, d, k_search, ef, qf, q_off.ctypes.data, stride, k,
                                   k_clusters, D.ctypes.data, I.ctypes.data, sw_scores.ctypes.data, sw_ids.ctypes.data,
                                   C.byref(st), tm.ctypes.data))
    return D, I, sw_scores, sw_ids, {f: getattr(st, f) for f, _ in SearchStats._fields_}, tm


def fasta_clean(data: bytes, device: int = 0) -> bytes:
    """extract_FASTA_sequence (src/utils/parse_inputs.cpp:174-220) on the device: file bytes in, cleaned genome out."""
    data = bytes(data)
    src = C.create_string_buffer(data, len(data)) if data else C.create_string_buffer(1)
    out = C.create_string_buffer(max(len(data), 1))
    ln = C.c_int64(0)
    _check(lib().drm_fasta_clean(C.addressof(src), len(data), device, C.addressof(out), len(data), C.byref(ln)))
    return out.raw[:ln.value]


class Reads:
    """format_fastq (src/utils/parse_inputs.cpp:843-950) on the device: the reads of a FASTQ file ('<' + sequence + '>')
    resident on every listed GPU."""

    def __init__(self, fastq: bytes, devices=None):
        arr, n = _devs(devices)
        data = bytes(fastq)
        src = C.create_string_buffer(data, len(data)) if data else C.create_string_buffer(1)
        h = C.c_void_p()
        _check(lib().drm_reads_open_fastq(C.addressof(src), len(data), arr, n, C.byref(h)))
        self._h = h
        info = np.zeros(4, dtype=np.int64)
        _check(lib().drm_reads_info(h, info.ctypes.data))
        self.n, self.nbytes, self.max_len, self.n_replicas = (int(v) for v in info)

    def close(self):
        if getattr(self, "_h", None):
            lib().drm_reads_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sequences(self):
        """Host copy: list of bytes, one per read (decorated like the reference's query_seqs)."""
        flat = C.create_string_buffer(max(self.nbytes, 1))
        off = np.zeros(self.n + 1, dtype=np.int64)
        _check(lib().drm_reads_copy_out(self._h, C.addressof(flat), off.ctypes.data))
        raw = flat.raw
        return [raw[off[i]:off[i + 1]] for i in range(self.n)]

    def device_ptrs(self, slot=0):
        f, o = C.c_void_p(), C.c_void_p()
        _check(lib().drm_reads_device(self._h, slot, C.byref(f), C.byref(o)))
        return f.value, o.value


def search_rerank_reads(index: Index, ref: Reference, mode, q, reads: Reads, stride, k, k_clusters, k_search, ef):
    """drm_search_rerank with the read sequences taken from a device-resident Reads handle."""
    n, d = q.shape
    D = np.empty((n, k_search), np.float32)
    I = np.empty((n, k_search), np.int64)
    sw_scores = np.empty((n, k), np.int32)
    sw_ids = np.empty((n, k), np.uint64)
    st = SearchStats()
    tm = np.zeros(2, dtype=np.float64)
    _check(lib().drm_search_rerank_reads(index._h, ref._h, mode, q.ctypes.data, n, d, k_search, ef, reads._h, stride, k, k_clusters,
                                         D.ctypes.data, I.ctypes.data, sw_scores.ctypes.data, sw_ids.ctypes.data, C.byref(st),
                                         tm.ctypes.data))
    return D, I, sw_scores, sw_ids, {f: getattr(st, f) for f, _ in SearchStats._fields_}, tm


def merge_topk_device(device, D_all_ptr, I_all_ptr, n_shards, n, k, D_out_ptr, I_out_ptr, stream=0):
    _check(lib().drm_merge_topk_device(device, D_all_ptr, I_all_ptr, n_shards, n, k, D_out_ptr, I_out_ptr, stream or None))


def build_index(x, out_path, M_pq=8, nbits=8, M_hnsw=16, efc=200, train=None, threads=0, seed=1234):
    """hnswpq_index's build_faiss_index (src/hnswpq/index.cpp:86-193) for .npy embeddings."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, d = x.shape
    if train is None:
        step = 2  # Config::Build::SAMPLE_RATE = 0.5, evenly strided (src/hnswpq/index.cpp:57-84,99)
        train = x[::step]
    train = np.ascontiguousarray(train, dtype=np.float32)
    _check(lib().drm_build_index(x.ctypes.data, n, d, train.ctypes.data, train.shape[0], M_pq, nbits, M_hnsw, efc, threads,
                                 seed, os.fsencode(out_path)))


BUILD_STAGES = ("upload", "pq_train", "encode", "graph", "export")


def build_index_gpu(x, out_path, M_pq=8, nbits=8, M_hnsw=16, efc=200, train=None, device=0, seed=1234):
    """build_faiss_index (src/hnswpq/index.cpp:86-193) on the GPU: PQ k-means, encoding and batched HNSW insertion
    (csrc/build_gpu.cu).  x: host array n x d.  Returns {stage: seconds}."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, d = x.shape
    if train is None:
        train = x[::2]  # Config::Build::SAMPLE_RATE = 0.5, evenly strided (src/hnswpq/index.cpp:57-84,99)
    train = np.ascontiguousarray(train, dtype=np.float32)
    tim = np.zeros(5, dtype=np.float64)
    _check(lib().drm_build_index_gpu(x.ctypes.data, n, d, train.ctypes.data, train.shape[0], M_pq, nbits, M_hnsw, efc, device,
                                     seed, os.fsencode(out_path), tim.ctypes.data))
    return dict(zip(BUILD_STAGES, tim.tolist()))


def build_index_gpu_device(x_dev_ptr, n, d, out_path, train_step=2, M_pq=8, nbits=8, M_hnsw=16, efc=200, device=0, seed=1234):
    """Same with the vectors already resident on `device` (x_dev_ptr: device address of n x d fp32); the PQ training
    sample is rows 0, train_step, 2*train_step, ...  Returns {stage: seconds}."""
    tim = np.zeros(5, dtype=np.float64)
    _check(lib().drm_build_index_gpu_device(x_dev_ptr, n, d, train_step, M_pq, nbits, M_hnsw, efc, device, seed,
                                            os.fsencode(out_path), tim.ctypes.data))
    return dict(zip(BUILD_STAGES, tim.tolist()))


# ---- embedder -------------------------------------------------------------------------------------
MAX_LEN = 123  # Config::Inference::MAX_LEN (includes/utils/config.hpp:21)


def tokenize(seqs, max_len=MAX_LEN):
    """Preprocessor::preprocessBatch (src/inference/preprocess.cpp:44-77): n x max_len vocabulary ids, 0-padded."""
    flat, off = flatten(seqs)
    out = np.zeros((len(off) - 1, max_len), dtype=np.uint16)
    _check(lib().drm_tokenize(flat, off.ctypes.data, len(off) - 1, max_len, out.ctypes.data))
    return out


def embedder_ir_info(xml_path, bin_path):
    """Host-side look at an OpenVINO IR pair: (vocabulary rows, element counts, sums) of emb, W1, R1, B1, W2, R2, B2."""
    info = np.zeros(8, dtype=np.int64)
    sums = np.zeros(7, dtype=np.float64)
    _check(lib().drm_embedder_ir_info(os.fsencode(xml_path), os.fsencode(bin_path), info.ctypes.data, sums.ctypes.data))
    return int(info[0]), info[1:].copy(), sums


class Vectorizer:
    """Vectorizer (src/inference/vectorize.cpp:4-141) on the GPU: tokeniser + token embedding + 2-layer bidirectional GRU.
    Open it on the reference's model files (model_path = the .xml; the .bin sits next to it) or on a dict of fp16 arrays
    {emb, W1, R1, B1, W2, R2, B2} in the IR's layout."""

    def __init__(self, model_path=None, weights=None, max_len=MAX_LEN, device=0):
        h = C.c_void_p()
        if weights is not None:
            a = {k: np.ascontiguousarray(weights[k], dtype=np.float16) for k in ("emb", "W1", "R1", "B1", "W2", "R2", "B2")}
            want = {"W1": (2, 192, 64), "R1": (2, 192, 64), "B1": (2, 256), "W2": (2, 192, 128), "R2": (2, 192, 64), "B2": (2, 256)}
            for k, shp in want.items():
                if a[k].shape != shp:
                    raise DrmError(-1, f"embedder weight {k} has shape {a[k].shape}, expected {shp}")
            if a["emb"].ndim != 2 or a["emb"].shape[1] != 64:
                raise DrmError(-1, "embedder weight emb must be [vocab, 64]")
            _check(lib().drm_embedder_open_raw(a["emb"].ctypes.data, a["emb"].shape[0], a["W1"].ctypes.data, a["R1"].ctypes.data,
                                               a["B1"].ctypes.data, a["W2"].ctypes.data, a["R2"].ctypes.data, a["B2"].ctypes.data,
                                               max_len, device, C.byref(h)))
        else:
            if max_len != MAX_LEN:
                raise DrmError(-1, "an IR model is opened with the reference's MAX_LEN")
            xml = os.fspath(model_path)
            _check(lib().drm_embedder_open_ir(os.fsencode(xml), os.fsencode(os.path.splitext(xml)[0] + ".bin"), device, C.byref(h)))
        self._h = h
        self.device = device
        self.max_len = max_len

    def close(self):
        if getattr(self, "_h", None):
            lib().drm_embedder_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def vectorize(self, seqs):
        """Vectorizer::vectorize: byte strings (decorated or not, exactly as given) -> float32 [n, 128]."""
        flat, off = flatten(seqs)
        n = len(off) - 1
        out = np.zeros((n, 128), dtype=np.float32)
        _check(lib().drm_embed(self._h, flat, off.ctypes.data, n, out.ctypes.data))
        return out

    def vectorize_reads(self, reads: "Reads", first=0, n=None):
        n = reads.n - first if n is None else n
        out = np.zeros((n, 128), dtype=np.float32)
        _check(lib().drm_embed_reads(self._h, reads._h, first, n, out.ctypes.data))
        return out

    def vectorize_windows(self, ref: "Reference", first_id, id_step, n):
        out = np.zeros((n, 128), dtype=np.float32)
        _check(lib().drm_embed_windows(self._h, ref._h, first_id, id_step, n, out.ctypes.data))
        return out

    def vectorize_device(self, flat_ptr, off_ptr, first, n, out_ptr, stream=0):
        _check(lib().drm_embed_device(self._h, flat_ptr, off_ptr, first, n, out_ptr, stream))

    def vectorize_windows_device(self, ref: "Reference", first_id, id_step, n, out_ptr, stream=0):
        _check(lib().drm_embed_windows_device(self._h, ref._h, first_id, id_step, n, out_ptr, stream))


# ---- L2 rerank stage (the reference's post_process_l2_* variants) -----------------------------------------------------
def l2_last_timing():
    """{plan_ms, embed_ms, rank_ms, other_ms, batches} of this thread's last drm_l2_rerank call."""
    t = np.zeros(5, dtype=np.float64)
    _check(lib().drm_l2_last_timing(t.ctypes.data))
    return dict(zip(("plan_ms", "embed_ms", "rank_ms", "other_ms", "batches"), t.tolist()))


def l2_plan_host(mode, neighbors, stride, k, k_clusters, bound, direct_ids=False):
    """The candidate bookkeeping of post_process_l2_static (mode 0) / _dynamic (mode 1) on the host:
    dict(unique, map, starts, k_out, passthrough, take).  direct_ids: the plan drm_l2_rerank makes when the window cache applies -
    `map` carries the window ids themselves and `unique` stays empty (no de-duplication pass)."""
    if direct_ids:
        mode = -1 - mode
    nb = np.ascontiguousarray(neighbors, dtype=np.int64)
    n, k_in = nb.shape
    info = np.zeros(5, dtype=np.int64)
    _check(lib().drm_l2_plan_host(mode, nb.ctypes.data, n, k_in, stride, k, k_clusters, bound, info.ctypes.data, None, None, None))
    unique = np.zeros(int(info[0]), dtype=np.uint64)
    mp = np.zeros(int(info[1]), dtype=np.uint32)
    starts = np.zeros(n + 1, dtype=np.int64)
    if not info[3]:
        _check(lib().drm_l2_plan_host(mode, nb.ctypes.data, n, k_in, stride, k, k_clusters, bound, info.ctypes.data,
                                      unique.ctypes.data, mp.ctypes.data, starts.ctypes.data))
    return dict(unique=unique, map=mp, starts=starts, k_out=int(info[2]), passthrough=bool(info[3]), take=int(info[4]))


def l2_select_host(dist, k):
    """std::partial_sort(idx, idx + k, idx + n, dist[a] < dist[b]) as the kernel runs it: the first k indices."""
    d = np.ascontiguousarray(dist, dtype=np.float32)
    idx = np.zeros(len(d), dtype=np.int64)
    _check(lib().drm_l2_select_host(d.ctypes.data, len(d), k, idx.ctypes.data))
    return idx[:k]


def _post_process_l2(mode, vectorizer, ref, neighbors, distances, query_embeddings, stride, k, k_clusters):
    nb = np.ascontiguousarray(neighbors, dtype=np.int64)
    ds = np.ascontiguousarray(distances, dtype=np.float32)
    qe = np.ascontiguousarray(query_embeddings, dtype=np.float32)
    n, k_in = nb.shape
    cap = n * max(k, k_clusters, 1)
    scores = np.zeros(cap, dtype=np.float32)
    ids = np.zeros(cap, dtype=np.uint64)
    cnt = C.c_int64(0)
    _check(lib().drm_l2_rerank(vectorizer._h, ref._h, mode, nb.ctypes.data, ds.ctypes.data, n, k_in, qe.ctypes.data, stride, k, k_clusters,
                               scores.ctypes.data, ids.ctypes.data, C.byref(cnt)))
    return scores[:cnt.value].copy(), ids[:cnt.value].copy()


def post_process_l2_dynamic(neighbors, distances, ref: "Reference", query_embeddings, vectorizer: "Vectorizer", stride, k, k_clusters=5):
    """post_process_l2_dynamic (src/utils/post_processor.cpp:552-748): (scores, dense ids), flat."""
    return _post_process_l2(1, vectorizer, ref, neighbors, distances, query_embeddings, stride, k, k_clusters)


def post_process_l2_static(neighbors, distances, ref: "Reference", query_embeddings, vectorizer: "Vectorizer", stride, k, k_clusters=5,
                           decorated=True):
    """post_process_l2_static (src/utils/post_processor.cpp:1023-1162): (scores, ids), k_clusters per query, flat.
    decorated=True: ref_seqs are '<' window '>' strings; False: the bare windows read_file(..., lookup_mode=true) returns, which
    is what the reference's main passes (src/main.cpp:190) - candidates are then embedded undecorated (device mode 2)."""
    return _post_process_l2(0 if decorated else 2, vectorizer, ref, neighbors, distances, query_embeddings, stride, k, k_clusters)


# ---- position-sharded index (BASELINE.json configs[4]) ------------------------------------------------
def _prefer_bundled_nccl():
    """libdrm_cuda.so dlopens NCCL at first use (an already loaded libnccl.so.2, else $DRM_NCCL_LIB, else the system one).  In a
    Python process that will ALSO import torch, the NCCL that torch was built against (the nvidia-nccl wheel) has to be the one
    in the process - torch fails to import on top of an older libnccl.so.2 - so point DRM_NCCL_LIB at the wheel's library when it
    exists and nothing else was asked for.  No torch import happens here."""
    if os.environ.get("DRM_NCCL_LIB"):
        return
    import importlib.util
    try:
        spec = importlib.util.find_spec("nvidia.nccl")
    except (ImportError, ValueError):
        spec = None
    for base in (list(spec.submodule_search_locations) if spec and spec.submodule_search_locations else []):
        cand = os.path.join(base, "lib", "libnccl.so.2")
        if os.path.exists(cand):
            os.environ["DRM_NCCL_LIB"] = cand
            return


class Shards:
    """One HNSW-PQ graph per GPU, windows cut by position (drm_shards_*).  Either all shards in this process
    (index_files / id_lo / devices lists: one host thread per GPU inside the calls) or this process's shard of a
    rank-per-process job (rank / world / unique_id as handed out by Shards.unique_id() on rank 0)."""

    def __init__(self, index_files, id_lo, devices, rank=None, world=None, unique_id=None):
        _prefer_bundled_nccl()
        h = C.c_void_p()
        if rank is None:
            files = (C.c_char_p * len(index_files))(*[os.fsencode(f) for f in index_files])
            lo = np.ascontiguousarray(id_lo, dtype=np.int64)
            arr, n = _devs(devices)
            assert n == len(index_files) == len(lo)
            _check(lib().drm_shards_open(files, lo.ctypes.data, arr, n, C.byref(h)))
        else:
            uid = C.create_string_buffer(bytes(unique_id) if unique_id else b"\0" * 128, 128)
            _check(lib().drm_shards_open_rank(os.fsencode(index_files), int(id_lo), int(devices), int(rank), int(world), C.addressof(uid), C.byref(h)))
        self._h = h
        info = np.zeros(9, dtype=np.int64)
        _check(lib().drm_shards_info(h, info.ctypes.data))
        (self.world, self.n_local, self.one_process, self.peer_ok, self.ntotal_local, self.device_bytes, self.rank, self.id_lo,
         self.nccl_version) = (int(v) for v in info)

    @staticmethod
    def unique_id() -> bytes:
        _prefer_bundled_nccl()
        buf = C.create_string_buffer(128)
        _check(lib().drm_shards_unique_id(C.addressof(buf)))
        return buf.raw

    def close(self):
        if getattr(self, "_h", None):
            lib().drm_shards_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def search_rerank(self, q, k_search, ef, exchange=XCHG_FUSED, chunk=0, ref=None, mode=1, reads=None, stride=1, k=None, k_clusters=None,
                      D=None, I=None, sw_scores=None, sw_ids=None):
        """Host buffers in, host buffers out.  reads = (flat bytes or address, int64 offsets) when ref is given.
        Returns (D, I, sw_scores | None, sw_ids | None, stats dict of local rank 0)."""
        q = np.ascontiguousarray(q, dtype=np.float32)
        n, d = q.shape
        D = np.full((n, k_search), np.nan, np.float32) if D is None else D
        I = np.full((n, k_search), -2, np.int64) if I is None else I
        st = ShardStats()
        if ref is not None:
            k = k_search if k is None else k
            k_clusters = k_search if k_clusters is None else k_clusters
            flat, off = reads
            keep = flat
            qf = flat if isinstance(flat, int) else C.cast(C.c_char_p(keep), C.c_void_p).value
            off = np.ascontiguousarray(off, dtype=np.int64)
            sw_scores = np.zeros((n, k), np.int32) if sw_scores is None else sw_scores
            sw_ids = np.zeros((n, k), np.uint64) if sw_ids is None else sw_ids
            _check(lib().drm_shards_search_rerank(self._h, ref._h, mode, q.ctypes.data, n, d, k_search, ef, qf, off.ctypes.data, stride, k, k_clusters,
                                                  exchange, chunk, D.ctypes.data, I.ctypes.data, sw_scores.ctypes.data, sw_ids.ctypes.data, C.byref(st)))
        else:
            _check(lib().drm_shards_search_rerank(self._h, None, 0, q.ctypes.data, n, d, k_search, ef, None, None, 1, k_search, k_search,
                                                  exchange, chunk, D.ctypes.data, I.ctypes.data, None, None, C.byref(st)))
        return D, I, sw_scores, sw_ids, {f: getattr(st, f) for f, _ in ShardStats._fields_}

    def step_device(self, q_ptr, n, d, k, ef, exchange=XCHG_FUSED, local_rank=0):
        """One chunk, device buffers (rank-per-process model): returns (D_ptr, I_ptr, slice_lo, slice_n, stream) - the merged
        lists of this rank's slice of the chunk, enqueued on the handle's stream."""
        Dp, Ip, st = C.c_void_p(), C.c_void_p(), C.c_void_p()
        lo, cnt = C.c_int64(0), C.c_int64(0)
        _check(lib().drm_shards_step_device(self._h, local_rank, q_ptr, n, d, k, ef, exchange, None, C.byref(Dp), C.byref(Ip), C.byref(lo), C.byref(cnt),
                                            C.byref(st)))
        return Dp.value, Ip.value, lo.value, cnt.value, st.value

    def stats(self, reset=False, local_rank=0):
        st = ShardStats()
        _check(lib().drm_shards_stats(self._h, local_rank, 1 if reset else 0, C.byref(st)))
        return {f: getattr(st, f) for f, _ in ShardStats._fields_}
