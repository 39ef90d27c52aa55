"""GPU parity tests for the embedder through the C-ABI: the CUDA path against the committed golden
vectors (reference tokeniser + torch.nn.GRU on the reference's weights) and against oracle/embed_oracle.py on seeded
inputs.  Floating point: the network runs in fp32 on fp16-exact weights; outputs lie in (-1, 1) and differ from the
oracle only by summation order and the exp/tanh implementations - the tolerance is 2e-4 absolute."""
import numpy as np
import pytest

import deepreadmapper_b200 as drm
from oracle import embed_oracle as EO

import embed_util as U

pytestmark = pytest.mark.gpu
TOL = 2e-4


@pytest.fixture(scope="module")
def gold():
    return U.load_golden()


@pytest.fixture(autouse=True, params=["fp32", "tc"])
def embed_form(request, monkeypatch):
    """Every test runs on both forms of the recurrences: the fp32 CUDA-core kernels and the tensor-core form (tcgen05 + TMEM,
    fp16 hi + lo activations on fp16-exact weights).  Same tolerance for both."""
    monkeypatch.setenv("DRM_EMBED_TC", "1" if request.param == "tc" else "0")
    return request.param


def _random_seqs(g, n, lo, hi, alphabet=b"ACGT", decorate=0.5):
    out = []
    for _ in range(n):
        s = bytes(g.choice(np.frombuffer(alphabet, np.uint8), int(g.integers(lo, hi + 1))))
        if g.random() < decorate:
            s = b"<" + s + b">"
        out.append(s)
    return out


def test_reference_weights_against_golden(gold):
    v = drm.Vectorizer(weights=gold["weights"])
    got = v.vectorize(gold["seqs"])
    assert got.shape == gold["embeddings"].shape
    assert float(np.abs(got - gold["embeddings"]).max()) < TOL
    # every golden tokeniser sequence too, against the oracle
    want = EO.embed(gold["tok_seqs"], gold["weights"])
    assert float(np.abs(v.vectorize(gold["tok_seqs"]) - want).max()) < TOL
    v.close()


def test_ragged_and_undefined_sequences_against_oracle():
    w = EO.random_weights(11)
    g = np.random.default_rng(12)
    seqs = [b"", b"A", b"AC", b"<A>", b"NN", b"ACGTNACGT" * 20, b"<ACGNNT>"]
    seqs += _random_seqs(g, 20, 2, 300) + _random_seqs(g, 18, 100, 160, alphabet=b"ACGTacgtN")
    assert len(seqs) % 16 != 0  # a partly filled last CTA
    v = drm.Vectorizer(weights=w)
    got = v.vectorize(seqs)
    want = EO.embed(seqs, w)
    assert float(np.abs(got - want).max()) < TOL
    assert v.vectorize([]).shape == (0, 128)
    one = v.vectorize(seqs[9:10])
    assert float(np.abs(one - want[9:10]).max()) < TOL
    v.close()


def test_more_sequences_than_one_pass(embed_form):
    import torch
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    n = (sms * 128 * 2 if embed_form == "tc" else sms * 16 * 4) + 37  # one full pass of the scratch + a short second one
    w = EO.random_weights(21)
    g = np.random.default_rng(22)
    seqs = _random_seqs(g, n, 100, 152, decorate=1.0)
    v = drm.Vectorizer(weights=w)
    got = v.vectorize(seqs)
    # sequences are independent: check the first and last CTAs of both passes and a random sample in between
    pick = np.unique(np.concatenate([np.arange(0, 48), np.arange(n - 37 - 48, n), g.integers(0, n, 96)]))
    want = EO.embed([seqs[i] for i in pick], w)
    assert float(np.abs(got[pick] - want).max()) < TOL
    assert np.isfinite(got).all() and float(np.abs(got).max()) <= 1.0
    v.close()


@pytest.mark.parametrize("max_len", [16, 2, 3, 5, 128])
def test_other_max_len(max_len):
    """Step counts around the tensor-core form's pipeline depths (layer 2 loads one step and prefetches four steps ahead), and the
    largest the kernels take."""
    w = EO.random_weights(31)
    g = np.random.default_rng(32)
    seqs = _random_seqs(g, 133, 2, 160)
    v = drm.Vectorizer(weights=w, max_len=max_len)
    assert float(np.abs(v.vectorize(seqs) - EO.embed(seqs, w, max_len=max_len)).max()) < TOL
    v.close()


def test_reads_handle_windows_and_ir_file(tmp_path):
    w = EO.random_weights(41)
    g = np.random.default_rng(42)
    v = drm.Vectorizer(weights=w)
    # FASTQ parsed on the device -> embedded in place (reads are held as '<' + sequence + '>')
    raw = _random_seqs(g, 40, 30, 150, decorate=0.0)
    fastq = b"".join(b"@r%d\n%s\n+\n%s\n" % (i, s, b"I" * len(s)) for i, s in enumerate(raw))
    reads = drm.Reads(fastq)
    want = EO.embed([b"<" + s + b">" for s in raw], w)
    assert float(np.abs(v.vectorize_reads(reads) - want).max()) < TOL
    assert float(np.abs(v.vectorize_reads(reads, first=7, n=12) - want[7:19]).max()) < TOL
    reads.close()
    # reference windows by find_sequence id (id / 2 = position, id % 2 = reverse complement), one id past the end
    genome = bytes(g.choice(np.frombuffer(b"ACGTN", np.uint8), 400, p=[0.245, 0.245, 0.245, 0.245, 0.02]))
    ref_len = 150
    ref = drm.Reference(genome=genome, ref_len=ref_len)
    comp = bytes.maketrans(b"ACGTN", b"TGCAN")
    n_pos = len(genome) - ref_len + 1
    ids = [3 + 5 * i for i in range(2 * n_pos // 5 + 1)]
    strings = []
    for i in ids:
        p = i // 2
        if p + ref_len > len(genome):
            strings.append(b"")
            continue
        win = genome[p:p + ref_len]
        strings.append(b"<" + (win.translate(comp)[::-1] if i % 2 else win) + b">")
    assert strings[-1] == b"" and strings[-2] != b""
    got = v.vectorize_windows(ref, 3, 5, len(ids))
    assert float(np.abs(got - EO.embed(strings, w)).max()) < TOL
    ref.close()
    # the same weights through an IR pair on disk
    xml = U.write_ir(str(tmp_path / "model"), w)
    v2 = drm.Vectorizer(model_path=xml)
    seqs = _random_seqs(g, 20, 100, 152, decorate=1.0)
    assert np.array_equal(v2.vectorize(seqs), v.vectorize(seqs))
    v2.close()
    v.close()


def test_cli_from_sequence_files_only(tmp_path):
    """hnswpq_index on a FASTA and pipeline on a FASTQ with no .npy beside either: both embed on the GPU through the
    model named by DRM_MODEL, like the reference's binaries do through OpenVINO (src/hnswpq/index.cpp:266-282,
    src/main.cpp:246-268)."""
    import os
    import subprocess
    from oracle import bindings as ob
    root = U.ROOT
    w = EO.random_weights(51)
    xml = U.write_ir(str(tmp_path / "model"), w)
    g = np.random.default_rng(52)
    genome = bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), 3000))
    ref_len = 150
    with open(tmp_path / "ref.fna", "wb") as f:
        f.write(b"> synthetic\n" + b"\n".join(genome[i:i + 70] for i in range(0, len(genome), 70)) + b"\n")
    comp = bytes.maketrans(b"ACGT", b"TGCA")
    reads = []
    for i in range(48):
        p = int(g.integers(0, len(genome) - ref_len))
        r = bytearray(genome[p:p + ref_len])
        for j in g.integers(0, ref_len, 3):
            r[j] = b"ACGT"[int(g.integers(0, 4))]
        r = bytes(r)
        reads.append(r.translate(comp)[::-1] if i % 2 else r)
    with open(tmp_path / "reads.fastq", "wb") as f:
        for i, r in enumerate(reads):
            f.write(b"@r%d\n%s\n+\n%s\n" % (i, r, b"I" * len(r)))
    env = dict(os.environ, DRM_MODEL=xml)
    r = subprocess.run([os.path.join(root, "build/bin/hnswpq_index"), "ref.fna", "idx1", "150", "1", "8", "8", "16", "64"], cwd=tmp_path,
                       capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr + r.stdout
    assert "Vectorizing" in r.stdout
    n_pos = len(genome) - ref_len + 1
    oix = ob.OracleIndex(str(tmp_path / "idx1" / "idx1.index"))
    assert oix.ntotal == 2 * n_pos and oix.d == 128
    # the vectors behind the codes are the windows in the reference's order (forward, reverse complement per position)
    v = drm.Vectorizer(model_path=xml)
    wins = []
    for p in range(n_pos):
        win = genome[p:p + ref_len]
        wins += [b"<" + win + b">", b"<" + win.translate(comp)[::-1] + b">"]
    E = v.vectorize(wins)
    cent, codes, _ = oix.arrays()
    dec = np.concatenate([cent[m][codes[:, m]] for m in range(cent.shape[0])], axis=1)
    right = float(((dec - E) ** 2).sum(1).mean())
    wrong = float(((dec - np.roll(E, 1, axis=0)) ** 2).sum(1).mean())
    assert right < 0.5 * wrong, (right, wrong)
    # pipeline without reads.fastq.npy: INFERENCE STEP on the GPU, then search + rerank; raw search output = the oracle's
    out = tmp_path / "out"
    r = subprocess.run([os.path.join(root, "build/bin/pipeline"), "idx1", "reads.fastq", "ref.fna", "64", "16", "16", str(out), "1"], cwd=tmp_path,
                       capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr + r.stdout
    assert "INFERENCE STEP" in r.stdout
    Q = v.vectorize([b"<" + x + b">" for x in reads])
    v.close()
    I, D = np.load(out / "indices.npy"), np.load(out / "distances.npy")
    D0, I0, _ = oix.search(Q, 16, 64, tie_rule=0)
    assert np.array_equal(I.astype(np.int64), I0) and np.array_equal(D.view(np.uint32), D0.view(np.uint32))
    S, SI = np.load(out / "sw_scores.npy"), np.load(out / "sw_indices.npy")
    rc, s0, i0, _ = ob.post_process_sw(1, I0, genome, len(genome), 150, ["<" + x.decode() + ">" for x in reads], 1, 16, 16)
    assert rc == 0 and np.array_equal(S, s0) and np.array_equal(SI, i0)


def test_prefetch_variant_is_bit_identical(monkeypatch):
    """DRM_EMBED_PREFETCH=1 (input projections loaded before a step's products; opt-in until it has been timed) does the
    same arithmetic in the same order as the default kernels."""
    w = EO.random_weights(61)
    g = np.random.default_rng(62)
    seqs = _random_seqs(g, 70, 2, 200)
    v = drm.Vectorizer(weights=w)
    monkeypatch.setenv("DRM_EMBED_TC", "0")  # both runs on the fp32 kernels, whatever form the fixture selected
    monkeypatch.setenv("DRM_EMBED_PREFETCH", "0")
    a = v.vectorize(seqs)
    monkeypatch.setenv("DRM_EMBED_PREFETCH", "1")
    b = v.vectorize(seqs)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert float(np.abs(b - EO.embed(seqs, w)).max()) < TOL
    v.close()


def test_tensor_core_form_against_fp32_kernels(monkeypatch, gold):
    """The tensor-core form against the fp32 kernels on the reference's own weights and on random weights: same inputs, both
    within TOL of the oracle, and the two forms within 5e-5 of each other (hi + lo activations carry 22 bits)."""
    g = np.random.default_rng(72)
    for w, seqs in ((gold["weights"], list(gold["seqs"]) + _random_seqs(g, 150, 100, 152, decorate=1.0)),
                    (EO.random_weights(71), _random_seqs(g, 300, 2, 200))):
        v = drm.Vectorizer(weights=w)
        monkeypatch.setenv("DRM_EMBED_TC", "0")
        a = v.vectorize(seqs)
        monkeypatch.setenv("DRM_EMBED_TC", "1")
        b = v.vectorize(seqs)
        b2 = v.vectorize(seqs)
        v.close()
        want = EO.embed(seqs, w)
        d_ab, d_bo = float(np.abs(a - b).max()), float(np.abs(b - want).max())
        print(f"tensor-core vs fp32 kernels: max abs diff {d_ab:.3g}; vs oracle {d_bo:.3g}")
        assert np.array_equal(b.view(np.uint32), b2.view(np.uint32))  # deterministic
        assert d_bo < TOL and d_ab < 5e-5


@pytest.mark.parametrize("tps", ["11", "12", "21", "24"])
def test_tensor_core_form_threads_per_sequence(monkeypatch, tps):
    """The gate phase with 1 / 2 / 4 threads per sequence (DRM_EMBED_TC_TPS = layer 1, layer 2; default 22): same products, same
    per-unit arithmetic, so the outputs are bit-identical to the default's."""
    w = EO.random_weights(81)
    g = np.random.default_rng(82)
    seqs = _random_seqs(g, 300, 2, 200)
    v = drm.Vectorizer(weights=w)
    monkeypatch.setenv("DRM_EMBED_TC", "1")
    monkeypatch.delenv("DRM_EMBED_TC_TPS", raising=False)
    a = v.vectorize(seqs)
    monkeypatch.setenv("DRM_EMBED_TC_TPS", tps)
    b = v.vectorize(seqs)
    v.close()
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert float(np.abs(b - EO.embed(seqs, w)).max()) < TOL
