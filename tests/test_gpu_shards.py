"""Position-sharded index through the C ABI (drm_shards_*, csrc/shards.cu): BASELINE.json configs[4] in miniature.
Parity target: the CPU oracle run the same sharded way (tests/shards_util.py) - ids and fp32 distance bits - and the
oracle's SW rerank of the merged labels.  The three exchange modes (NCCL all-gather, NCCL all-to-all, fused peer-memory
writes from the search kernel) must return identical bits.  Multi-GPU cases skip on a one-GPU box; they are run with
`gpurun --gpus 2/8` and their logs kept under profiles/."""
import os
import subprocess
import sys

import numpy as np
import pytest

import shards_util as SU

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def drm():
    import deepreadmapper_b200 as d
    assert d.lib().drm_device_count() >= 1, "no CUDA device"
    return d


@pytest.fixture(scope="module")
def world_data(drm, tmp_path_factory):
    from deepreadmapper_b200 import sharded, synth
    g = synth.random_genome(30000, 11)
    reads, truth = synth.simulate_reads(g, 700, seed=5)
    Q = synth.embed_reads(reads)
    out = {}
    for W in (1, 2, 3):
        d = str(tmp_path_factory.mktemp(f"shards{W}"))
        out[W] = SU.build_shards(drm, synth, sharded, g, W, d)
    return g, reads, truth, Q, out


def _rerank_oracle(ob, g, I, reads, k):
    rc, s0, i0, _ = ob.post_process_sw(1, I, g, len(g), 150, [bytes(r) for r in reads], 1, k, I.shape[1])
    assert rc == 0
    return s0, i0


def test_one_shard_handle_equals_plain_index(drm, ob, world_data):
    """W = 1: no exchange; the handle must return what drm_search + drm_sw_rerank return, with labels offset by id_lo."""
    g, reads, truth, Q, shards = world_data
    paths, los, _ = shards[3]           # take the MIDDLE shard of the 3-way split: id_lo > 0
    path, lo = paths[1], los[1]
    k, ef = 32, 64
    sh = drm.Shards([path], [lo], [0])
    assert (sh.world, sh.n_local, sh.id_lo) == (1, 1, lo)
    ref = drm.Reference(genome=g, ref_len=150, devices=[0])
    flat, off = drm.api.flatten([bytes(r) for r in reads])
    for chunk in (0, 256, 300):
        D, I, s, ids, st = sh.search_rerank(Q, k, ef, chunk=chunk, ref=ref, mode=1, reads=(flat, off), k=16)
        D0, I0 = SU.oracle_sharded_search(ob, [path], [lo], Q, k, ef)
        assert np.array_equal(I, I0) and np.array_equal(D.view(np.uint32), D0.view(np.uint32)), chunk
        s0, i0 = _rerank_oracle(ob, g, I0, reads, 16)
        assert np.array_equal(s, s0) and np.array_equal(ids, i0), chunk
        assert st["nhops0"] > 0 and st["steps"] > 0
    sh.close()
    ref.close()


def test_shards_errors(drm, world_data, tmp_path):
    g, reads, truth, Q, shards = world_data
    paths, los, _ = shards[1]
    with pytest.raises(drm.DrmError) as e:
        drm.Shards([str(tmp_path / "missing.index")], [0], [0])
    assert "Index file does not exist" in str(e.value)
    with pytest.raises(drm.DrmError):
        drm.Shards([paths[0], paths[0]], [0, 0], [0, 0])       # one distinct GPU per shard
    sh = drm.Shards(paths, los, [0])
    with pytest.raises(drm.DrmError) as e:
        sh.search_rerank(np.zeros((0, 128), np.float32), 8, 16)
    assert str(e.value) == "Query data is empty"
    with pytest.raises(drm.DrmError):
        sh.search_rerank(np.zeros((4, 64), np.float32), 8, 16)
    sh.close()


@pytest.mark.parametrize("W", [2, 3])
def test_shards_one_process_all_exchange_modes(drm, ob, world_data, W):
    """W GPUs driven by one process (what `pipeline` does under DRM_SHARD=position)."""
    if drm.lib().drm_device_count() < W:
        pytest.skip(f"needs {W} GPUs")
    g, reads, truth, Q, shards = world_data
    paths, los, _ = shards[W]
    k, ef = 32, 64
    D0, I0 = SU.oracle_sharded_search(ob, paths, los, Q, k, ef)
    s0, i0 = _rerank_oracle(ob, g, I0, reads, 16)
    sh = drm.Shards(paths, los, list(range(W)))
    ref = drm.Reference(genome=g, ref_len=150, devices=list(range(W)))
    flat, off = drm.api.flatten([bytes(r) for r in reads])
    for exchange in (drm.XCHG_FUSED, drm.XCHG_ALLTOALL, drm.XCHG_ALLGATHER, drm.XCHG_FUSED):
        for chunk in (0, 250):
            D, I, s, ids, st = sh.search_rerank(Q, k, ef, exchange=exchange, chunk=chunk, ref=ref, mode=1, reads=(flat, off), k=16)
            assert np.array_equal(I, I0) and np.array_equal(D.view(np.uint32), D0.view(np.uint32)), (exchange, chunk)
            assert np.array_equal(s, s0) and np.array_equal(ids, i0), (exchange, chunk)
    assert float(np.mean(np.any((I0 >> 1) == (truth[:, None] >> 1), axis=1))) > 0.5
    sh.close()
    ref.close()


@pytest.mark.parametrize("W", [2])
def test_shards_rank_per_process(drm, ob, world_data, W, tmp_path):
    """One process per GPU (torchrun-style): tests/shards_rank_check.py run as W ranks, the ncclUniqueId passed through a file."""
    if drm.lib().drm_device_count() < W:
        pytest.skip(f"needs {W} GPUs")
    g, reads, truth, Q, shards = world_data
    paths, los, _ = shards[W]
    np.save(tmp_path / "Q.npy", Q)
    k, ef = 32, 64
    D0, I0 = SU.oracle_sharded_search(ob, paths, los, Q, k, ef)
    np.save(tmp_path / "D0.npy", D0)
    np.save(tmp_path / "I0.npy", I0)
    procs = []
    for r in range(W):
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "shards_rank_check.py"), str(r), str(W), paths[r], str(los[r]),
                                       str(tmp_path), str(k), str(ef)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=600)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and "rank ok" in o, (r, o[-3000:])
