"""Run-time FAISS probe.  If a real FAISS is importable where the tests run, the
restatement (oracle/hnswpq_oracle.cpp, tie_rule 0), its IHNp reader and - on a GPU box - the CUDA kernel and drm_index_open
are pinned against a FAISS-BUILT file and real IndexHNSWPQ::search.  If not, the probe's finding is printed and the pin is
skipped: the search half stays "parity unpinned" (DESIGN.md §2)."""
import numpy as np
import pytest

from deepreadmapper_b200 import synth
from oracle import faiss_probe, parity


def test_probe_reports():
    p = faiss_probe.probe()
    print("[faiss probe]", p)
    assert set(p) >= {"found", "python", "shared_lib", "how"}
    assert faiss_probe.search_kind(p) in ("faiss", "restatement")


def _pin(tmp_path, ob, drm=None):
    p = faiss_probe.probe()
    if not p["python"]:
        pytest.skip(f"FAISS is not importable here ({p.get('python_error')}; shared lib: {p['shared_lib']}): search parity stays unpinned")
    g = synth.random_genome(1000, 42)  # BASELINE.json configs[0] size: 1 702 vectors
    E = synth.embed_windows(g)
    reads, _ = synth.simulate_reads(g, 150)
    Q = synth.embed_reads(reads)
    path = str(tmp_path / "faiss_built.index")
    Df, If = faiss_probe.pin_against_faiss(E, Q, 128, 128, path, efc=200)
    oix = ob.OracleIndex(path)                      # the restatement's reader on a file FAISS wrote
    assert oix.validate() == 0 and oix.ntotal == len(E) and (oix.M_pq, oix.nbits, oix.M_hnsw) == (8, 8, 16)
    D0, I0, _ = oix.search(Q, 128, 128, tie_rule=0)
    rep = parity.tie_rule_report(Df, If, D0, I0)
    print("[faiss pin] real FAISS vs restatement:", rep)
    assert rep["rows_violating"] == 0 and rep["overlap"] >= 0.99, rep
    if drm is not None:
        ix = drm.Index(path)                        # drm_index_open on the FAISS-written file
        D1, I1 = ix.search(Q, 128, 128)
        ix.close()
        rep1 = parity.tie_rule_report(Df, If, D1, I1)
        print("[faiss pin] real FAISS vs CUDA kernel:", rep1)
        assert rep1["rows_violating"] == 0 and rep1["overlap"] >= 0.99, rep1


def test_restatement_against_real_faiss(tmp_path, ob):
    _pin(tmp_path, ob)


@pytest.mark.gpu
def test_kernel_against_real_faiss(tmp_path, ob):
    import deepreadmapper_b200 as drm
    _pin(tmp_path, ob, drm)
