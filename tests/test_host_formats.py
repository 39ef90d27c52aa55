"""CPU tests of the host C++ side (deepreadmapper_b200/host): file formats and CLIs, compared with what the
UNMODIFIED reference code (oracle/_ref) produces for the same inputs.  No GPU is needed: nothing here computes on the
hot path; the CLIs are checked for their format handling and for failing loudly when there is no device."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "build", "bin")


@pytest.fixture(scope="module")
def tool():
    if not os.path.exists(os.path.join(BIN, "formats_tool")):
        subprocess.check_call(["make", "-s", "-C", ROOT, "host"])
    return os.path.join(BIN, "formats_tool")


@pytest.fixture(scope="module")
def R(ob):
    if not ob.have_ref():
        pytest.skip("oracle/_ref (the compiled reference) is not present")
    return ob.ref()


def _write_fasta(path, contigs, width=70, lower=False):
    with open(path, "w") as f:
        for i, s in enumerate(contigs):
            f.write(f">contig{i} Fake DNA sequence ACGT\n")
            s = s.lower() if lower else s
            for p in range(0, len(s), width):
                f.write(s[p:p + width] + "\n")


def test_save_results_files_are_byte_identical_to_the_reference(tool, R, tmp_path):
    rng = np.random.default_rng(1)
    for n, k_in, k in [(7, 16, 16), (3, 20, 5), (1, 1, 1), (129, 128, 128)]:
        I = rng.integers(0, 10 ** 7, (n, k_in)).astype(np.int64)
        I[rng.random((n, k_in)) < 0.1] = -1  # unfilled labels -> 2^64-1 in indices.npy
        D = rng.random((n, k_in)).astype(np.float32)
        I.tofile(tmp_path / "I.bin")
        D.tofile(tmp_path / "D.bin")
        mine = [str(tmp_path / "mi.npy"), str(tmp_path / "md.npy")]
        ref = [str(tmp_path / "ri.npy"), str(tmp_path / "rd.npy")]
        subprocess.check_call([tool, "save_results", str(tmp_path / "I.bin"), str(tmp_path / "D.bin"), str(n), str(k_in), str(k)] + mine)
        assert R.ref_save_results(I.reshape(-1), D.reshape(-1), n, k_in, k, ref[0].encode(), ref[1].encode()) == 0
        for a, b in zip(mine, ref):
            assert open(a, "rb").read() == open(b, "rb").read(), (n, k_in, k, a)
        got = np.load(mine[0])
        assert got.dtype == np.dtype("<u8") and got.shape == (n, k) and np.array_equal(got, I[:, :k].astype(np.uint64))
        assert np.array_equal(np.load(mine[1]), D[:, :k])


def test_sam_file_is_byte_identical_to_the_reference(tool, R, tmp_path):
    """write_sam (src/utils/utils.cpp:336-403): FLAG / POS from the dense id, pseudo MAPQ and CIGAR, '<' '>' stripped from
    SEQ, FASTQ names as QNAME; fewer results than n x k cut the tail; an empty result list leaves the header only."""
    if not hasattr(R, "ref_write_sam"):
        pytest.skip("oracle/_ref predates ref_write_sam (rebuild with /root/reference present)")
    rng = np.random.default_rng(5)
    fq = tmp_path / "reads.fastq"
    with open(fq, "w") as f:
        for i, ln in enumerate([150, 150, 97, 1, 151, 150, 33]):
            f.write("@read%d extra words\n%s\n+\n%s\n" % (i, "".join(rng.choice(list("ACGTN"), ln)), "I" * ln))
    n = 7
    for k, n_results in [(4, 28), (1, 7), (5, 23), (3, 0), (2, 14)]:
        ids = rng.integers(0, 2 * 10 ** 7, n * k).astype(np.uint64)
        ids.tofile(tmp_path / "ids.bin")
        mine, ref = str(tmp_path / "mine.sam"), str(tmp_path / "ref.sam")
        subprocess.check_call([tool, "sam", str(fq), str(tmp_path / "ids.bin"), str(n_results), "ref", "150", str(k), mine], stdout=subprocess.DEVNULL)
        assert R.ref_write_sam(str(fq).encode(), ids.ctypes.data, len(ids), n_results, b"ref", 150, k, ref.encode()) == 0
        a, b = open(mine, "rb").read(), open(ref, "rb").read()
        assert a == b, (k, n_results)
        assert a.count(b"\n") == 2 + min(n_results, n * k)


def test_npy_reader_accepts_numpy_and_cnpy_headers(tool, tmp_path):
    a = np.arange(24, dtype=np.float32).reshape(6, 4)
    np.save(tmp_path / "a.npy", a)
    out = subprocess.check_output([tool, "npyinfo", str(tmp_path / "a.npy")]).decode()
    assert out.startswith("f4 fortran=0 shape=6,4, bytes=96")
    np.save(tmp_path / "b.npy", np.arange(5, dtype=np.uint64))
    assert subprocess.check_output([tool, "npyinfo", str(tmp_path / "b.npy")]).decode().startswith("u8 fortran=0 shape=5, bytes=40")


def test_fasta_windows_and_labels_match_the_reference(tool, R, fixtures, tmp_path):
    g = fixtures["genome"]
    cases = [("single.fna", [g], False), ("multi.fna", [g[:400], g[400:520], g[520:]], False), ("lower.fa", [g[:300] + "NNxx" + g[300:]], True)]
    for name, contigs, lower in cases:
        path = str(tmp_path / name)
        _write_fasta(path, contigs, lower=lower)
        for ref_len, stride, lookup in [(150, 1, 1), (150, 8, 0), (97, 3, 1)]:
            n, nb = C.c_int64(0), C.c_int64(0)
            assert R.ref_preprocess_fasta(path.encode(), ref_len, stride, lookup, None, None, None, C.byref(n), C.byref(nb)) == 0
            flat = C.create_string_buffer(max(nb.value, 1))
            off = np.zeros(n.value + 1, dtype=np.int64)
            lab = np.zeros(max(n.value, 1), dtype=np.uint64)
            assert R.ref_preprocess_fasta(path.encode(), ref_len, stride, lookup, flat, off.ctypes.data, lab.ctypes.data, C.byref(n), C.byref(nb)) == 0
            want = [flat.raw[off[i]:off[i + 1]].decode("latin1") for i in range(n.value)]
            subprocess.check_call([tool, "fasta", path, str(ref_len), str(stride), str(lookup), str(tmp_path / "w.txt"), str(tmp_path / "l.txt")])
            got = [l for l in open(tmp_path / "w.txt", encoding="latin1").read().split("\n") if l]
            got_l = [int(x) for x in open(tmp_path / "l.txt").read().split()]
            assert got == want, (name, ref_len, stride, lookup)
            assert got_l == [int(v) for v in lab[:n.value]]
        # the cleaned genome of the dynamic mode
        cap = 1 << 16
        buf = C.create_string_buffer(cap)
        ln = R.ref_extract_fasta_sequence(path.encode(), buf, cap)
        subprocess.check_call([tool, "genome", path, str(tmp_path / "g.txt")])
        assert open(tmp_path / "g.txt").read() == buf.raw[:ln].decode()
    # config 1 of BASELINE.json: the single-contig fixture gives the window table the goldens were made from
    subprocess.check_call([tool, "fasta", str(tmp_path / "single.fna"), "150", "1", "1", str(tmp_path / "w.txt"), str(tmp_path / "l.txt")])
    assert [l for l in open(tmp_path / "w.txt").read().split("\n") if l] == fixtures["windows"]


def test_fastq_sequences_and_ids_match_the_reference(tool, R, fixtures, tmp_path):
    path = str(tmp_path / "reads.fastq")
    with open(path, "w") as f:
        for i, (rid, seq) in enumerate(zip(fixtures["read_ids"][:30], fixtures["reads"][:30])):
            s = seq.strip("<>")
            if i == 4:
                s = s.lower()
            sep = [" extra", "\tx", "/1", ""][i % 4]
            f.write(f"@{rid}{sep}\n{s}\n+\n{'I' * (len(s) - (i % 3))}\n")  # short quality lines, as in tests/test_data.fastq
    n, sb, ib = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    assert R.ref_preprocess_fastq(path.encode(), None, None, None, None, C.byref(n), C.byref(sb), C.byref(ib)) == 0
    sflat, iflat = C.create_string_buffer(sb.value + 1), C.create_string_buffer(ib.value + 1)
    soff, ioff = np.zeros(n.value + 1, np.int64), np.zeros(n.value + 1, np.int64)
    assert R.ref_preprocess_fastq(path.encode(), sflat, soff.ctypes.data, iflat, ioff.ctypes.data, C.byref(n), C.byref(sb), C.byref(ib)) == 0
    want_s = [sflat.raw[soff[i]:soff[i + 1]].decode() for i in range(n.value)]
    want_i = [iflat.raw[ioff[i]:ioff[i + 1]].decode() for i in range(n.value)]
    subprocess.check_call([tool, "fastq", path, str(tmp_path / "s.txt"), str(tmp_path / "i.txt")])
    assert open(tmp_path / "s.txt").read().split("\n")[:-1] == want_s
    assert open(tmp_path / "i.txt").read().split("\n")[:-1] == want_i
    assert want_s[0].startswith("<") and want_s[0].endswith(">")


def test_config_txt_round_trip_with_the_reference(tool, R, tmp_path):
    folder = str(tmp_path / "idx")
    assert R.ref_save_config(folder.encode(), 8, 150, 1160376, 128, 16, 200, 8, 8, b"idx/idx.index") == 0
    out = str(tmp_path / "idx2")
    subprocess.check_call([tool, "config", folder + "/config.txt", out])  # our load_config -> our save_config
    for key, val in [("stride", 8), ("ref_len", 150), ("n_vects", 1160376), ("dim", 128), ("M_hnsw", 16), ("EFC", 200), ("M_pq", 8), ("nbits", 8)]:
        assert R.ref_load_config_size((out + "/config.txt").encode(), key.encode()) == val
    lines = dict(l.split(": ", 1) for l in open(out + "/config.txt").read().strip().split("\n"))
    assert lines["index_type"] == "HNSWPQ" and lines["index_file"] == "idx/idx.index"


def test_hnswpq_index_cli_writes_the_reference_layout(R, ob, tmp_path):
    """hnswpq_index <emb.npy> <prefix> <ref_len> ... -> <prefix>/<prefix>.index + config.txt.  The CLI builds on the GPU and has
    no silent fallback: without a device it fails loudly unless DRM_BUILD=cpu opts into the host OpenMP builder (used here);
    DRM_SHARDS=<W> writes W position shards and the config keys `pipeline` reads."""
    from deepreadmapper_b200 import synth
    g = synth.random_genome(3000, 5)
    np.save(tmp_path / "ref.npy", synth.embed_windows(g))
    exe = os.path.join(BIN, "hnswpq_index")
    import deepreadmapper_b200 as drm
    env_cpu = dict(os.environ, DRM_BUILD="cpu")
    if drm.lib().drm_device_count() == 0:
        env_none = {k: v for k, v in os.environ.items() if k != "DRM_BUILD"}
        r = subprocess.run([exe, "ref.npy", "myidx", "150", "1", "8", "8", "16", "40"], cwd=tmp_path, capture_output=True, text=True, env=env_none)
        assert r.returncode == 1 and "no CUDA device" in r.stderr and "DRM_BUILD=cpu" in r.stderr
    r = subprocess.run([exe, "ref.npy", "myidx", "150", "1", "8", "8", "16", "40"], cwd=tmp_path, capture_output=True, text=True, env=env_cpu)
    assert r.returncode == 0, r.stderr
    # position shards: 3 files with local labels + n_shards / shard_<g>_id_lo in config.txt
    r = subprocess.run([exe, "ref.npy", "shidx", "150", "1", "8", "8", "16", "40"], cwd=tmp_path, capture_output=True, text=True, env=dict(env_cpu, DRM_SHARDS="3"))
    assert r.returncode == 0, r.stderr
    scfg = str(tmp_path / "shidx" / "config.txt").encode()
    assert R.ref_load_config_size(scfg, b"n_shards") == 3 and R.ref_load_config_size(scfg, b"shard_0_id_lo") == 0
    n_all, tot = 2 * (3000 - 149), 0
    for g_ in range(3):
        lo = R.ref_load_config_size(scfg, b"shard_%d_id_lo" % g_)
        assert lo % 2 == 0 and lo == 2 * (((n_all + 1) // 2) * g_ // 3)
        so = ob.OracleIndex(str(tmp_path / "shidx" / f"shidx.shard{g_}.index"))
        assert so.validate() == 0
        tot += so.ntotal
    assert tot == n_all
    cfg = str(tmp_path / "myidx" / "config.txt").encode()
    assert R.ref_load_config_size(cfg, b"ref_len") == 150 and R.ref_load_config_size(cfg, b"stride") == 1
    assert R.ref_load_config_size(cfg, b"n_vects") == 2 * (3000 - 149) and R.ref_load_config_size(cfg, b"EFC") == 40
    oix = ob.OracleIndex(str(tmp_path / "myidx" / "myidx.index"))
    assert oix.validate() == 0
    # usage and error behaviour of the reference CLI
    assert subprocess.run([exe], capture_output=True).returncode == 1
    r = subprocess.run([exe, "missing.fna", "x", "150"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 1 and r.stderr.startswith("Error: Embeddings not found")


def test_pipeline_cli_argument_and_file_errors(tmp_path):
    exe = os.path.join(BIN, "pipeline")
    r = subprocess.run([exe, "a", "b"], capture_output=True, text=True)
    assert r.returncode == 1 and r.stderr.startswith("Usage:")
    r = subprocess.run([exe, str(tmp_path / "noidx"), "q.npy", "ref.fna"], capture_output=True, text=True)
    assert r.returncode == 1 and r.stderr.strip() == f"Error: Config file does not exist: {tmp_path}/noidx/config.txt"
    os.makedirs(tmp_path / "idx")
    open(tmp_path / "idx" / "config.txt", "w").write("ref_len: 150\nstride: 1\n")
    np.save(tmp_path / "q.npy", np.zeros((4, 128), np.float32))
    r = subprocess.run([exe, str(tmp_path / "idx"), str(tmp_path / "q.npy"), "ref.fna"], capture_output=True, text=True)
    assert r.returncode == 1 and r.stderr.strip() == f"Error: Index file does not exist: {tmp_path}/idx/idx.index"
    np.save(tmp_path / "q3.npy", np.zeros((4, 2, 2), np.float32))
    r = subprocess.run([exe, str(tmp_path / "idx"), str(tmp_path / "q3.npy"), "ref.fna"], capture_output=True, text=True)
    assert r.returncode == 1 and r.stderr.strip() == "Error: Error: Expected 2D array in .npy file"
