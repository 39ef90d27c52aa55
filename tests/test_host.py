"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol include/drm_b200.h
declares, the index builder writes FAISS-format files the oracle's independent reader accepts, the compute
entry points fail loudly without a GPU, and the oracle's two tie rules agree up to PQ-distance ties."""
import os
import re
import struct

import numpy as np
import pytest

import deepreadmapper_b200 as drm
from deepreadmapper_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "drm_b200.h")).read()
    declared = set(re.findall(r"\b(drm_[a-z0-9_]+)\s*\(", hdr))
    assert declared and declared == set(drm.api.EXPORTS)
    L = drm.lib()
    for name in declared:
        assert getattr(L, name) is not None
    assert b"sm_100a" in L.drm_version()


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "deepreadmapper_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".hpp")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in src and "from oracle" not in src and "liboracle" not in src


@pytest.fixture(scope="module")
def built(tmp_path_factory):
    g = synth.random_genome(8000, 42)
    E = synth.embed_windows(g)
    p = str(tmp_path_factory.mktemp("idx") / "t.index")
    # one thread: the host builder's OpenMP insertion order is not deterministic, and the tie-rule overlap below sits
    # within a few 1e-3 of its bound
    drm.build_index(E, p, M_pq=8, nbits=8, M_hnsw=16, efc=64, seed=5, threads=1)
    return g, E, p


def test_builder_writes_faiss_layout(built, ob):
    g, E, p = built
    with open(p, "rb") as f:
        raw = f.read(64)
    assert raw[:4] == b"IHNp"
    d, ntotal = struct.unpack_from("<iq", raw, 4)
    assert d == 128 and ntotal == len(E) == 2 * (8000 - 150 + 1)
    ix = ob.OracleIndex(p)
    assert ix.validate() == 0
    assert (ix.d, ix.ntotal, ix.M_pq, ix.nbits, ix.M_hnsw, ix.efConstruction) == (128, len(E), 8, 8, 16, 64)
    assert ix.max_level >= 1


def test_builder_graph_finds_planted_windows(built, ob):
    g, E, p = built
    reads, truth = synth.simulate_reads(g, 300, sub_rate=0.01, indel_rate=0.0)
    Q = synth.embed_reads(reads)
    ix = ob.OracleIndex(p)
    D, I, st = ix.search(Q, 128, 128, tie_rule=0)
    De, Ie = ix.adc_scan(Q, 128)
    recall = np.mean([len(set(I[i].tolist()) & set(Ie[i].tolist())) / 128 for i in range(len(Q))])
    assert recall > 0.5, recall
    near = np.mean([(np.abs((I[i] >> 1) - (truth[i] >> 1)) <= 2).any() for i in range(len(Q))])
    assert near > 0.8, near


def test_oracle_tie_rules_agree_up_to_ties(built, ob):
    g, E, p = built
    reads, _ = synth.simulate_reads(g, 400)
    Q = synth.embed_reads(reads)
    ix = ob.OracleIndex(p)
    D0, I0, s0 = ix.search(Q, 64, 128, tie_rule=0)
    D1, I1, s1 = ix.search(Q, 64, 128, tie_rule=1)
    overlap = np.mean([len(set(I0[i].tolist()) & set(I1[i].tolist())) / 64 for i in range(len(Q))])
    assert overlap >= 0.99, overlap
    # canonical output is ordered by (distance, id) and self-consistent
    assert (np.diff(D1, axis=1) >= 0).all()
    for i in range(0, len(Q), 25):
        assert np.array_equal(ix.adc_distances(Q[i], I1[i]).view(np.uint32), D1[i].view(np.uint32))


def test_compute_calls_fail_loudly_without_gpu(built):
    if drm.lib().drm_device_count() > 0:
        pytest.skip("a GPU is present")
    g, E, p = built
    with pytest.raises(drm.DrmError) as e:
        drm.Index(p)
    assert e.value.code == -3 and "no CPU fallback" in str(e.value)
    with pytest.raises(drm.DrmError) as e:
        drm.sw_pairs(["ACGT"], ["ACGT"])
    assert e.value.code == -3
    with pytest.raises(drm.DrmError):
        drm.Reference(genome=g, ref_len=150)
    with pytest.raises(drm.DrmError) as e:
        drm.build_index_gpu(E[:200], p + ".gpu", efc=16)
    assert e.value.code == -3 and "no CPU fallback" in str(e.value)
    for call in (lambda: drm.fasta_clean(b">h\nACGT\n"), lambda: drm.Reads(b"@r\nACGT\n+\nIIII\n"),
                 lambda: drm.Reference(fasta=b">h\nACGT\n", ref_len=2)):
        with pytest.raises(drm.DrmError) as e:
            call()
        assert e.value.code == -3 and "no CPU fallback" in str(e.value)


def test_index_open_errors_without_touching_gpu(tmp_path):
    with pytest.raises(drm.DrmError) as e:
        drm.Index(str(tmp_path / "nope.index"))
    assert e.value.code == -2 and "Index file does not exist" in str(e.value)
    bad = tmp_path / "bad.index"
    bad.write_bytes(b"IxF2" + b"\0" * 100)
    with pytest.raises(drm.DrmError) as e:
        drm.Index(str(bad))
    assert e.value.code == -2


def test_synth_window_ids_match_reference_layout(fixtures):
    """Row 2p / 2p+1 of embed_windows are the forward / reverse-complement window at base p (tests/test_data_ref.txt order)."""
    g = fixtures["genome"].encode()
    E = synth.embed_windows(g)
    assert E.shape == (fixtures["n_windows"], 128)
    w = fixtures["windows"]
    for wid in (0, 1, 560, 561, 1701):
        e = synth.embed_reads(np.frombuffer(w[wid].encode(), dtype=np.uint8)[None, :])
        assert np.allclose(e[0], E[wid], atol=1e-4)
    assert synth.reverse_complement(w[10].encode()) == w[11].encode()
