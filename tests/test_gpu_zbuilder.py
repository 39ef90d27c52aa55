"""GPU index builder (csrc/build_gpu.cu; SURVEY.md section 8(f) rank 1 - the hnswpq_index half, src/hnswpq/index.cpp:86-193).
There is no bit-level oracle for graph construction (FAISS's own build is not reproducible under OpenMP, and neither is a
lock-ordered GPU build), so the bar is: the product quantizer equals the host builder's for the same seed, the file is a
valid FAISS IHNp index, and searching it recalls the exact-ADC neighbours at least as well as the host-built graph does."""
import numpy as np
import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


@pytest.fixture(scope="module")
def drm():
    import deepreadmapper_b200 as d
    assert d.lib().drm_device_count() >= 1, "no CUDA device"
    return d


@pytest.fixture(scope="module")
def data():
    from deepreadmapper_b200 import synth
    g = synth.random_genome(40000, 11)
    E = synth.embed_windows(g)
    reads, truth = synth.simulate_reads(g, 1500)
    return g, E, synth.embed_reads(reads), truth


@pytest.fixture(scope="module")
def both(tmp_path_factory, drm, data):
    g, E, Q, truth = data
    d = tmp_path_factory.mktemp("bld")
    pc, pg = str(d / "cpu.index"), str(d / "gpu.index")
    drm.build_index(E, pc, efc=100, seed=9)
    tim = drm.build_index_gpu(E, pg, efc=100, seed=9)
    return pc, pg, tim


def _recall(ob, path, Q, k, ef):
    ix = ob.OracleIndex(path)
    _, I, _ = ix.search(Q, k, ef, tie_rule=1)
    _, Ie = ix.adc_scan(Q, k)
    return float(np.mean([len(set(I[i].tolist()) & set(Ie[i].tolist())) / k for i in range(len(Q))]))


def test_product_quantizer_matches_host_builder(ob, both):
    pc, pg, tim = both
    cc, kc, _ = ob.OracleIndex(pc).arrays()
    cg, kg, _ = ob.OracleIndex(pg).arrays()
    assert np.allclose(cc, cg, rtol=1e-5, atol=1e-7), float(np.abs(cc - cg).max())
    assert np.mean(kc == kg) > 0.999
    assert set(tim) == {"upload", "pq_train", "encode", "graph", "export"} and tim["graph"] > 0


def test_graph_is_valid_faiss_index_with_host_builder_quality(drm, ob, data, both):
    g, E, Q, truth = data
    pc, pg, _ = both
    oc, og = ob.OracleIndex(pc), ob.OracleIndex(pg)
    assert og.validate() == 0
    assert (og.d, og.ntotal, og.M_pq, og.nbits, og.M_hnsw, og.efConstruction) == (128, len(E), 8, 8, 16, 100)
    assert og.max_level == oc.max_level and og.entry_point == oc.entry_point  # same seed -> same levels and order
    dc, dg = oc.arrays()[2], og.arrays()[2]
    assert dg["empty_rows0"] == 0.0
    assert 0.5 * dc["mean_degree0"] <= dg["mean_degree0"] <= 2.0 * dc["mean_degree0"], (dc, dg)
    for k, ef in ((10, 32), (64, 128)):
        rc, rg = _recall(ob, pc, Q, k, ef), _recall(ob, pg, Q, k, ef)
        assert rg >= 0.9 * rc - 0.02, (k, ef, rc, rg)
    # the search kernel on the GPU-built file: bit-exact against the oracle reading the same file
    D0, I0, _ = og.search(Q, 64, 128, tie_rule=0)
    ix = drm.Index(pg)
    D1, I1 = ix.search(Q, 64, 128)
    ix.close()
    assert np.array_equal(I0, I1) and np.array_equal(D0.view(np.uint32), D1.view(np.uint32))


def test_device_resident_build_same_quantizer(drm, ob, data, both, tmp_path):
    import torch
    g, E, Q, truth = data
    x = torch.from_numpy(E).cuda()
    p = str(tmp_path / "dev.index")
    drm.build_index_gpu_device(x.data_ptr(), x.shape[0], x.shape[1], p, train_step=2, efc=100, seed=9)
    od = ob.OracleIndex(p)
    assert od.validate() == 0
    c0, k0, _ = ob.OracleIndex(both[1]).arrays()
    c1, k1, _ = od.arrays()
    assert np.array_equal(c0, c1) and np.array_equal(k0, k1)  # the quantizer is deterministic
    assert _recall(ob, p, Q[:500], 10, 64) >= 0.9 * _recall(ob, both[0], Q[:500], 10, 64) - 0.03


@pytest.mark.parametrize("n,M_pq,M_hnsw,efc", [(1, 8, 16, 40), (5, 8, 16, 40), (700, 8, 4, 8), (3000, 16, 8, 300), (3000, 8, 16, 1)])
def test_small_and_odd_shapes(drm, ob, data, tmp_path, n, M_pq, M_hnsw, efc):
    E = data[1][:n]
    p = str(tmp_path / "s.index")
    drm.build_index_gpu(E, p, M_pq=M_pq, M_hnsw=M_hnsw, efc=efc, seed=4)
    ix = ob.OracleIndex(p)
    assert ix.validate() == 0
    assert (ix.ntotal, ix.M_pq, ix.M_hnsw) == (n, M_pq, M_hnsw)
    if n >= 700 and efc >= 8:
        k = 5
        _, I, _ = ix.search(E[:200], k, 64, tie_rule=1)
        _, Ie = ix.adc_scan(E[:200], k)
        rec = np.mean([len(set(I[i].tolist()) & set(Ie[i].tolist())) / k for i in range(200)])
        assert rec > 0.5, rec


def test_unsupported_shapes_are_refused(drm, data, tmp_path):
    E = data[1][:500]
    for kw in (dict(M_hnsw=32), dict(M_pq=4), dict(nbits=4)):
        with pytest.raises(drm.DrmError) as e:
            drm.build_index_gpu(E, str(tmp_path / "x.index"), **kw)
        assert e.value.code in (-1, -4)
