"""drm-b200: B200-native HNSW-PQ search + Smith-Waterman rerank (DeepReadMapper's search-time hot path).

The product is the C-ABI shared library ``deepreadmapper_b200/lib/libdrm_cuda.so`` (include/drm_b200.h) and
the C++ host binaries built on it.  This Python package is the thin ctypes mirror of the reference's
call sites used by tests/ and bench.py.  It never imports ``oracle`` and has no CPU compute path:
without the built library, or without a CUDA device, the compute calls raise.
"""
from .api import (  # noqa: F401
    DrmError, Index, Reads, Reference, Shards, XCHG_ALLGATHER, XCHG_ALLTOALL, XCHG_FUSED, Vectorizer, embedder_ir_info, tokenize, build_index, build_index_gpu, build_index_gpu_device, calc_sw_score, faiss_search, lib, lib_path, merge_topk_device,
    post_process_l2_dynamic, post_process_l2_static, l2_plan_host, l2_select_host, post_process_sw_dynamic, post_process_sw_static, fasta_clean, search_rerank, search_rerank_reads, sw_pairs, sw_pairs_device_noalloc, sw_reranker,
)
