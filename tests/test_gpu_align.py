"""SW traceback kernel (csrc/align_kernel.cu; SURVEY.md section 8(f) rank 3: the "real CIGAR" the reference's write_sam leaves
as a TODO, src/utils/utils.cpp:381-384).  The reference has no traceback, so the oracle (oracle/sw_oracle.c: orc_sw_align) is
a definition; it is pinned to the reference through the score: every alignment's score equals calc_sw_score and re-scoring
its operations gives the score back (tests/test_oracle_sw.py).  Here the kernel must equal that oracle field for field."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]

COMP = bytes.maketrans(b"ACGTN", b"TGCAN")


@pytest.fixture(scope="module")
def drm():
    import deepreadmapper_b200 as d
    assert d.lib().drm_device_count() >= 1, "no CUDA device"
    return d


def window(g: bytes, wid: int, ref_len: int) -> bytes:
    """find_sequence (src/utils/post_processor.cpp:47-66)"""
    pos, rc = wid >> 1, wid & 1
    if pos + ref_len > len(g):
        return b""
    w = g[pos:pos + ref_len]
    return w.translate(COMP)[::-1] if rc else w


def check(ob, got, g, ids, reads, ref_len, strip=0, max_ops=64):
    n, per = ids.shape
    for i in range(n):
        rd = bytes(reads[i])
        if strip and len(rd) > 2 * strip:
            rd = rd[strip:-strip]
        for s in range(per):
            want = ob.sw_align(window(g, int(ids[i, s]), ref_len), rd, max_ops)
            for k in ("score", "ref_beg", "ref_end", "q_beg", "q_end", "n_ops", "nm"):
                assert int(got[k][i, s]) == want[k], (i, s, k, int(got[k][i, s]), want)
            if want["n_ops"] > max_ops:
                continue  # truncated: only the run count is specified
            ops = [(int(w) >> 4, "MID"[int(w) & 15]) for w in got["ops"][i, s, :want["n_ops"]]]
            assert ops == want["ops"], (i, s, ops, want)


def test_alignments_match_the_oracle(drm, ob):
    from deepreadmapper_b200 import synth
    g = synth.random_genome(6000, 13)
    reads, truth = synth.simulate_reads(g, 300, sub_rate=0.05, indel_rate=0.02, seed=3)
    rng = np.random.default_rng(1)
    nwin = 2 * (len(g) - 149)
    ids = np.stack([truth, np.clip(truth + 2 * rng.integers(-6, 7, len(truth)), 0, nwin - 1), truth ^ 1,
                    rng.integers(0, nwin, len(truth))], axis=1).astype(np.uint64)
    ref = drm.Reference(genome=g, ref_len=150)
    seqs = [bytes(r) for r in reads]
    got = ref.align(ids, seqs, max_ops=160)
    check(ob, got, g, ids, seqs, 150, max_ops=160)
    assert (got["n_ops"] <= 160).all() and (got["n_ops"][:, 3] > 8).any()  # random windows give long meandering paths
    # score == calc_sw_score (the pinned SW path), and the planted window aligns (almost) end to end
    flat_ids = ids.reshape(-1)
    a = [window(g, int(w), 150) for w in flat_ids]
    b = [seqs[i // 4] for i in range(len(flat_ids))]
    assert np.array_equal(got["score"].reshape(-1), drm.sw_pairs(a, b))
    assert np.median(got["ref_end"][:, 0] - got["ref_beg"][:, 0]) >= 140
    # '<' '>' decoration stripped like write_sam does: same alignments
    got2 = ref.align(ids, [b"<" + s + b">" for s in seqs], strip=1, max_ops=160)
    for k in got:
        assert np.array_equal(got[k], got2[k]), k
    ref.close()


def test_alignment_edge_cases(drm, ob):
    g = b"ACGTTGCAAGGCTTAACCGGTTACGATCGATCGGATCCATGCAAATTTGGGCCCNNNNACGTACGTAGCTAGCTAGGATC" * 3
    ref_len = 40
    ref = drm.Reference(genome=g, ref_len=ref_len)
    nwin = 2 * (len(g) - ref_len + 1)
    reads = [g[10:50], g[10:50].translate(COMP)[::-1], b"A", b"", b"TTTTTTTTTT", g[5:25] + b"GG" + g[25:60], g[0:20] + g[23:70],
             b"NNNNACGT", (g + g)[:256], b"acgtacgt"]
    ids = np.array([[20, 21, 0, nwin + 10]] * len(reads), dtype=np.uint64)  # last id overruns the genome -> no alignment
    got = ref.align(ids, reads)
    check(ob, got, g, ids, reads, ref_len)
    assert (got["score"][:, 3] == 0).all() and (got["n_ops"][:, 3] == 0).all()
    assert got["score"][0, 0] == 40 and got["ops"][0, 0, 0] == (40 << 4) and got["score"][1, 1] == 40
    # a tiny operation buffer reports the true run count (> max_ops = truncated)
    t = ref.align(ids[5:7], reads[5:7], max_ops=1)
    assert (t["n_ops"][:, 0] >= 2).all()
    with pytest.raises(drm.DrmError):
        ref.align(ids[:1], [b"A" * 300])
    ref.close()
    # window-table form (ref_seqs[id])
    wins = [window(g, w, ref_len) for w in range(nwin)]
    tab = drm.Reference(windows=wins, ref_len=ref_len)
    got_t = tab.align(ids[:8, :3], reads[:8])
    for k in ("score", "ref_beg", "ref_end", "q_beg", "q_end", "n_ops", "nm"):
        assert np.array_equal(got_t[k], got[k][:8, :3]), k
    tab.close()


def test_pipeline_writes_sam_with_real_cigars(drm, tmp_path):
    """DRM_SAM=2 through the binaries: error-free reads from known positions come back as POS = position + 1, CIGAR 150M,
    NM 0, flag 0 / 16, and reverse-strand lines carry the reverse-complemented read."""
    from deepreadmapper_b200 import synth
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    g = synth.random_genome(5000, 77)
    open(tmp_path / "ref.fna", "wb").write(b"> synthetic\n" + b"\n".join(g[i:i + 80] for i in range(0, len(g), 80)) + b"\n")
    np.save(tmp_path / "ref.fna.npy", synth.embed_windows(g))
    reads, truth = synth.simulate_reads(g, 40, sub_rate=0.0, indel_rate=0.0, seed=2)
    with open(tmp_path / "reads.fastq", "wb") as f:
        for i, r in enumerate(reads):
            f.write(b"@read%d extra\n" % i + bytes(r) + b"\n+\n" + b"I" * 150 + b"\n")
    np.save(tmp_path / "reads.fastq.npy", synth.embed_reads(reads))
    r = subprocess.run([os.path.join(root, "build/bin/hnswpq_index"), "ref.fna", "idx", "150", "1", "8", "8", "16", "64"], cwd=tmp_path,
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = tmp_path / "out"
    r = subprocess.run([os.path.join(root, "build/bin/pipeline"), "idx", "reads.fastq", "ref.fna", "128", "64", "64", str(out), "1"], cwd=tmp_path,
                       capture_output=True, text=True, env=dict(os.environ, DRM_SAM="2"))
    assert r.returncode == 0, r.stderr + r.stdout
    lines = open(out / "alignments.sam").read().split("\n")
    assert lines[0] == "@HD\tVN:1.0\tSO:unsorted" and lines[1] == f"@SQ\tSN:ref\tLN:{len(g)}"
    body = [l.split("\t") for l in lines[2:] if l]
    assert len(body) == 2 * len(reads)
    S = np.load(out / "sw_scores.npy")
    found = 0
    for i in range(len(reads)):
        prim, sec = body[2 * i], body[2 * i + 1]
        assert prim[0] == f"read{i}" and int(prim[1]) & 256 == 0 and int(sec[1]) & 256 == 256
        assert f"AS:i:{S[i, 0]}" in prim
        if S[i, 0] == 150:  # the search found the planted window: the line must be the exact placement
            found += 1
            pos, rc = int(truth[i]) >> 1, int(truth[i]) & 1
            assert int(prim[3]) == pos + 1 and prim[5] == "150M" and "NM:i:0" in prim and (int(prim[1]) & 16 == 16) == bool(rc)
            assert prim[9].encode() == g[pos:pos + 150]  # forward-strand sequence whatever the read's strand
    assert found >= len(reads) // 2
