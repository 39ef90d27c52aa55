"""Size-independent properties of the hot path at BASELINE.json's FULL sizes (configs[1]: 9 283 006 windows, 1 000 000 reads,
EF = K = 128), where the oracle is too slow to be the checker:

  search    rows ascending in distance, ids unique per row and inside the index, padding only at the tail; the same call
            twice gives the same bits (the persistent kernel hands queries out through an atomic counter: order must not matter);
  rerank    scores descending per row (sw_reranker's order, src/utils/reranker.cpp:35-48) and inside [0, min(read, window)];
            the returned ids of a read are a PERMUTATION of its candidates (dense index, k = all candidates); idempotent;
  stage     the one-call host-buffer stage (drm_search_rerank: chunked, three streams, copies overlapped) returns, bit for
            bit, what the two device calls return - a checksum of every row, not a sample.

The index is the bench's (built once per box by the GPU builder and cached; seconds)."""
import argparse
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.fixture(scope="module")
def full():
    import torch

    import bench
    import deepreadmapper_b200 as drm
    args = argparse.Namespace(genome_bases=bench.GENOME_BASES, efc=200, builder="gpu", cache=os.environ.get("DRM_CACHE", "/tmp/drm_b200_cache"),
                              workload="dense")
    os.makedirs(args.cache, exist_ok=True)
    torch.cuda.set_device(0)
    genome = bench.make_genome(args)
    path, _ = bench.ensure_index(args, genome, "cuda:0")
    ix = drm.Index(path, devices=[0])
    ref = drm.Reference(genome=genome, ref_len=bench.REF_LEN, devices=[0])
    n = 1_000_000
    reads, truth, Q = bench.make_reads(genome, n, seed=7)
    yield dict(torch=torch, drm=drm, bench=bench, ix=ix, ref=ref, n=n, reads=reads, Q=Q, ntotal=2 * (len(genome) - bench.REF_LEN + 1))
    ix.close()
    ref.close()


def _device_stage(f):
    """search + rerank as the two device calls of the bench's device arm; returns the four result tensors (on the device)."""
    torch, K, EF, n = f["torch"], f["bench"].K, f["bench"].EF, f["n"]
    if "dev" not in f:
        f["dev"] = dict(Q=torch.from_numpy(f["Q"]).cuda(), reads=torch.from_numpy(f["reads"].reshape(-1)).cuda(),
                        off=torch.from_numpy(np.arange(n + 1, dtype=np.int64) * f["bench"].REF_LEN).cuda())
    d = f["dev"]
    D = torch.empty((n, K), dtype=torch.float32, device="cuda")
    I = torch.empty((n, K), dtype=torch.int64, device="cuda")
    S = torch.empty((n, K), dtype=torch.int32, device="cuda")
    W = torch.empty((n, K), dtype=torch.int64, device="cuda")
    sp = torch.cuda.current_stream().cuda_stream
    f["ix"].search_device(d["Q"].data_ptr(), n, K, EF, D.data_ptr(), I.data_ptr(), 0, sp)
    f["ref"].rerank_device(1, I.data_ptr(), n, K, d["reads"].data_ptr(), d["off"].data_ptr(), f["bench"].REF_LEN, 1, K, K, S.data_ptr(), W.data_ptr(),
                           0, 0, 0, sp)
    torch.cuda.synchronize()
    return D, I, S, W


def test_search_properties_at_full_size(full):
    torch = full["torch"]
    D, I, S, W = _device_stage(full)
    full["first"] = (D, I, S, W)
    n, K = I.shape
    pad = I < 0
    assert bool(((I >= 0) & (I < full["ntotal"]) | pad).all()), "a label outside the index"
    assert bool((pad[:, 1:] | ~pad[:, :-1]).all()), "padding before a result"       # once a row pads, it pads to the end
    assert bool((D[:, 1:] >= D[:, :-1]).all()), "a row is not ascending in distance"
    srt = torch.sort(I, dim=1).values
    dup = (srt[:, 1:] == srt[:, :-1]) & (srt[:, 1:] >= 0)
    assert not bool(dup.any()), "a vertex twice in one row"
    assert int(pad.sum()) == 0, "EF = K = 128 on a 9.28 M-vertex graph fills every row"


def test_rerank_properties_at_full_size(full):
    torch = full["torch"]
    D, I, S, W = full.get("first") or _device_stage(full)
    assert bool((S[:, 1:] <= S[:, :-1]).all()), "scores not descending"
    assert int(S.min()) >= 0 and int(S.max()) <= full["bench"].REF_LEN, "a score outside [0, min(read, window)]"
    assert bool((torch.sort(W, dim=1).values == torch.sort(I, dim=1).values).all()), "the reranked ids are not a permutation of the candidates"
    # most reads are simulated from the genome with 3 % substitutions: their best window must score far above a random pair's ~20
    assert float((S[:, 0] >= 100).float().mean()) > 0.05


def test_stage_is_deterministic_and_one_call_equals_two_calls_at_full_size(full):
    torch, drm, K, EF, n = full["torch"], full["drm"], full["bench"].K, full["bench"].EF, full["n"]
    first = full.get("first") or _device_stage(full)
    again = _device_stage(full)
    for a, b, name in zip(first, again, ("D", "I", "scores", "ids")):
        assert bool((a.view(torch.int32) == b.view(torch.int32)).all()) if a.dtype == torch.float32 else bool((a == b).all()), f"{name} differs between two runs"
    # the host-buffer stage: chunks, three streams, copies overlapped with kernels
    q_off = np.arange(n + 1, dtype=np.int64) * full["bench"].REF_LEN
    hreads = np.ascontiguousarray(full["reads"].reshape(-1))
    hD, hI, hS, hW, st, tm = drm.search_rerank(full["ix"], full["ref"], 1, full["Q"], hreads.ctypes.data, q_off, 1, K, K, K, EF)
    D, I, S, W = (t.cpu().numpy() for t in first)
    assert np.array_equal(hI, I) and np.array_equal(hD.view(np.uint32), D.view(np.uint32)), "search: host stage != device calls"
    assert np.array_equal(hS, S) and np.array_equal(hW.astype(np.int64), W), "rerank: host stage != device calls"
