"""Input side on the device (csrc/input_kernels.cu): FASTA -> cleaned genome and
FASTQ -> read sequences, bit-exact against the UNMODIFIED reference parsers (oracle/_ref: extract_FASTA_sequence,
preprocess_fastq -> format_fastq) and against a plain Python restatement of the two loops."""
import ctypes as C

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


@pytest.fixture(scope="module")
def drm():
    import deepreadmapper_b200 as d
    assert d.lib().drm_device_count() >= 1, "no CUDA device"
    return d


def py_extract_fasta(data: bytes) -> bytes:
    """src/utils/parse_inputs.cpp:174-220"""
    i = data.find(b"\n")
    body = data[i + 1:] if i >= 0 else b""
    keep = set(b"ACGTN")
    return bytes(c for c in body.upper() if c in keep)


def py_format_fastq(data: bytes):
    """src/utils/parse_inputs.cpp:843-950 (sequences only)"""
    seqs, i, line, n = [], 0, 0, len(data)
    while i < n:
        j = data.find(b"\n", i)
        if j < 0:
            j = n
        if line % 4 == 1:
            seqs.append(b"<" + data[i:j] + b">")
        i = j + 1
        line += 1
    return seqs


def ref_extract_fasta(ob, tmp_path, data):
    p = tmp_path / "x.fa"
    p.write_bytes(data)
    cap = len(data) + 16
    buf = C.create_string_buffer(cap)
    ln = ob.ref().ref_extract_fasta_sequence(str(p).encode(), buf, cap)
    assert ln >= 0
    return buf.raw[:ln]


def ref_format_fastq(ob, tmp_path, data):
    R = ob.ref()
    p = tmp_path / "x.fastq"
    p.write_bytes(data)
    n, sb, ib = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    assert R.ref_preprocess_fastq(str(p).encode(), None, None, None, None, C.byref(n), C.byref(sb), C.byref(ib)) == 0
    ids = max(n.value, (data.count(b"\n") + 4) // 4) + 2
    sflat, iflat = C.create_string_buffer(sb.value + 1), C.create_string_buffer(ib.value + 1)
    soff, ioff = np.zeros(n.value + 1, np.int64), np.zeros(ids + 1, np.int64)
    assert R.ref_preprocess_fastq(str(p).encode(), sflat, soff.ctypes.data, iflat, ioff.ctypes.data, C.byref(n), C.byref(sb), C.byref(ib)) == 0
    return [sflat.raw[soff[i]:soff[i + 1]] for i in range(n.value)]


def _fasta_cases():
    rng = np.random.default_rng(3)
    big = bytes(rng.choice(np.frombuffer(b"ACGTacgtNnRYKM\n\r \t>", dtype=np.uint8), size=3_000_017))
    return {
        "plain": b">chr1 test\nACGTACGTNN\nGGCC\n",
        "lower_iupac_crlf": b">h\r\nacgtnRYKM\r\nACGT\r\n",
        "second_header_letters_leak": b">one\nACGT\n>contig2 GATTACA tag\nTTTT\n",
        "no_trailing_newline": b">h\nACGT\nAC",
        "header_only": b">only a header\n",
        "no_newline_at_all": b">ACGTACGT",
        "empty": b"",
        "newline_first": b"\nACGT",
        "big_random": b">hdr with ACGT\n" + big,
    }


@pytest.mark.parametrize("name", list(_fasta_cases()))
def test_fasta_clean_matches_reference(drm, ob, tmp_path, name):
    data = _fasta_cases()[name]
    got = drm.fasta_clean(data)
    assert got == py_extract_fasta(data), name
    if ob.have_ref() and data:
        assert got == ref_extract_fasta(ob, tmp_path, data), name


def _fastq_cases():
    rng = np.random.default_rng(5)
    recs = []
    for i in range(60_000):
        L = int(rng.integers(0, 260)) if i % 97 == 0 else 150
        s = bytes(rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), size=L))
        recs.append(b"@read%d/1 extra\n" % i + s + b"\n+\n" + b"I" * L + b"\n")
    big = b"".join(recs)
    return {
        "plain": b"@r1\nACGT\n+\nIIII\n@r2 x\nGGCCA\n+\nIIIII\n",
        "no_trailing_newline": b"@r1\nACGT\n+\nIIII\n@r2\nGG\n+\nII",
        "sequence_is_last_line_without_newline": b"@r1\nACGT\n+\nIIII\n@r2\nGGTT",
        "header_only_last_record": b"@r1\nACGT\n+\nIIII\n@r2\n",
        "crlf_kept": b"@r1\r\nACGT\r\n+\r\nIIII\r\n",
        "empty_sequence_and_blank_lines": b"@r1\n\n+\n\n@r2\nAC\n+\nII\n\n\n",
        "empty": b"",
        "only_newlines": b"\n\n\n\n\n\n",
        "big": big,
    }


@pytest.mark.parametrize("name", list(_fastq_cases()))
def test_fastq_reads_match_reference(drm, ob, tmp_path, name):
    data = _fastq_cases()[name]
    want = py_format_fastq(data)
    rd = drm.Reads(data)
    got = rd.sequences()
    assert rd.n == len(want) and got == want, name
    assert rd.max_len == max([len(s) for s in want], default=0)
    rd.close()
    if ob.have_ref() and data:
        assert want == ref_format_fastq(ob, tmp_path, data), name


def test_reference_from_fasta_and_reads_from_fastq_feed_the_stage(drm, ob, tmp_path):
    """pipeline wiring: Reference(fasta=...) == Reference(genome=cleaned), and drm_search_rerank_reads (sequences resident on
    the device) == drm_search_rerank (host sequences), bit for bit."""
    from deepreadmapper_b200 import synth
    g = synth.random_genome(30000, 21)
    fasta = b">chrT synthetic\n" + b"\n".join(g[i:i + 70] for i in range(0, len(g), 70)).lower() + b"\n"
    assert drm.fasta_clean(fasta) == g
    reads, truth = synth.simulate_reads(g, 2500)
    fastq = b"".join(b"@q%d\n" % i + bytes(r) + b"\n+\n" + b"I" * len(r) + b"\n" for i, r in enumerate(reads))
    path = str(tmp_path / "t.index")
    drm.build_index_gpu(synth.embed_windows(g), path, efc=64, seed=3)
    Q = synth.embed_reads(reads)
    ix = drm.Index(path)
    ref_a = drm.Reference(genome=g, ref_len=150)
    ref_b = drm.Reference(fasta=fasta, ref_len=150)
    assert ref_b.length == len(g)
    rd = drm.Reads(fastq)
    assert rd.n == len(reads)
    seqs = [b"<" + bytes(r) + b">" for r in reads]  # what format_fastq hands to the SW stage
    flat, off = drm.api.flatten(seqs)
    D1, I1, s1, id1, _, _ = drm.search_rerank(ix, ref_a, 1, Q, bytes(flat), off, 1, 32, 64, 64, 128)
    D2, I2, s2, id2, st, _ = drm.search_rerank_reads(ix, ref_b, 1, Q, rd, 1, 32, 64, 64, 128)
    for a, b in ((D1, D2), (I1, I2), (s1, s2), (id1, id2)):
        assert np.array_equal(a, b)
    if drm.lib().drm_device_count() >= 2:  # the same with everything replicated on two GPUs (reads sharded inside the call)
        ix2, ref2, rd2 = drm.Index(path, devices=[0, 1]), drm.Reference(fasta=fasta, ref_len=150, devices=[0, 1]), drm.Reads(fastq, devices=[0, 1])
        assert rd2.n_replicas == 2
        D3, I3, s3, id3, _, _ = drm.search_rerank_reads(ix2, ref2, 1, Q, rd2, 1, 32, 64, 64, 128)
        for a, b in ((D1, D3), (I1, I3), (s1, s3), (id1, id3)):
            assert np.array_equal(a, b)
        for h in (rd2, ref2, ix2):
            h.close()
    # and both equal the oracle on the undecorated reads (the decoration never changes a score)
    rc, s0, i0, _ = ob.post_process_sw(1, I1, g, len(g), 150, [bytes(r) for r in reads], 1, 32, 64)
    assert rc == 0 and np.array_equal(s0, s1) and np.array_equal(i0, id1)
    for h in (rd, ref_a, ref_b, ix):
        h.close()
