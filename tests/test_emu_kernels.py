"""CPU tests that run the SOURCE of two CUDA kernels files on host threads (tools/cpu_emu: one OS thread per CUDA thread,
C++20 barriers for __syncthreads / __syncwarp, AddressSanitizer on) and compare with the oracle:

  * csrc/embed_kernel.cu - all three kernels, every source mode (byte strings, windows by id progression, windows by id
    list, bare windows) - after the hardware run of round 1 the file gained the id-list / bare-window source;
  * csrc/l2_kernel.cu    - the L2 rerank kernel, which has not run on hardware yet.

This checks indexing, phase ordering and arithmetic of the kernel code; it is not a substitute for the -m gpu tests
(nothing here exercises the device, its limits or its math library)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import embed_oracle as EO
from oracle import l2_oracle as L2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 2e-4
# The emulation runs one OS thread per CUDA thread: the complete set below (both library builds, every kernel family plain and
# under AddressSanitizer) takes about 20 minutes on 8 cores.  The default CPU suite therefore runs the LIGHT set - the embedder
# and L2 kernels, the search kernels and the merge / traceback / input kernels once each on the plain build - and
# DRM_EMU_FULL=1 adds the AddressSanitizer passes and the two end-to-end L2 stage checks (last full run:
# profiles/r2_pytest_cpu_full.log).
FULL = os.environ.get("DRM_EMU_FULL") == "1"
full_only = pytest.mark.skipif(not FULL, reason="heavy host-thread emulation pass: set DRM_EMU_FULL=1 (see profiles/r2_pytest_cpu_full.log)")


@pytest.fixture(scope="module")
def emu():
    sys.path.insert(0, os.path.join(ROOT, "tools", "cpu_emu"))
    import emu_build
    # AddressSanitizer's runtime has to be loaded first, so the sanitised library runs in a child interpreter (below);
    # the plain build is loaded here
    lib = C.CDLL(emu_build.build(sanitize=False))
    lib.emu_last_error.restype = C.c_char_p
    return lib


def _classes(genome: bytes) -> np.ndarray:
    t = np.full(256, 7, np.uint8)
    for i, c in enumerate(b"ACGT"):
        t[c] = i
    t[ord("N")] = 4
    return t[np.frombuffer(genome, np.uint8)]


def _embed(lib, w, max_len=123, seqs=None, cls=None, form=0, ref_len=150, first_id=0, id_step=1, ids=None, decorate=1, n=None):
    a = {k: np.ascontiguousarray(w[k], np.float16) for k in w}
    vp = lambda x: x.ctypes.data_as(C.c_void_p)  # noqa: E731
    if seqs is not None:
        flat = b"".join(seqs) + b"\0"
        off = np.zeros(len(seqs) + 1, np.int64); off[1:] = np.cumsum([len(s) for s in seqs])
        n = len(seqs)
        mode, flat_p, off_p, cls_p, ref_n = 0, C.c_char_p(flat), vp(off), None, 0
    else:
        mode, flat_p, off_p, cls_p, ref_n = 1, None, None, vp(cls), len(cls) if form == 0 else len(cls) // ref_len
    ids_p = None
    if ids is not None:
        ids = np.ascontiguousarray(ids, np.uint64); ids_p = vp(ids); n = len(ids)
    out = np.zeros((n, 128), np.float32)
    rc = lib.emu_embed(vp(a["emb"]), a["emb"].shape[0], vp(a["W1"]), vp(a["R1"]), vp(a["B1"]), vp(a["W2"]), vp(a["R2"]), vp(a["B2"]),
                       max_len, mode, flat_p, off_p, cls_p, form, C.c_int64(ref_n), ref_len, C.c_uint64(first_id), C.c_uint64(id_step), ids_p,
                       decorate, C.c_int64(0), C.c_int64(n), vp(out))
    assert rc == 0, lib.emu_last_error()
    return out


def _window(genome, dense_id, ref_len):
    comp = bytes.maketrans(b"ACGTN", b"TGCAN")
    p = dense_id // 2
    if p + ref_len > len(genome):
        return None
    w = genome[p:p + ref_len]
    return w.translate(comp)[::-1] if dense_id % 2 else w


@pytest.fixture(params=["0", "1"], ids=["default", "prefetch"])
def prefetch(request, monkeypatch):
    """Both code paths of gru_kernel: the default one and the DRM_EMBED_PREFETCH=1 variant (opt-in until measured)."""
    monkeypatch.setenv("DRM_EMBED_PREFETCH", request.param)
    return request.param


def test_embed_kernels_byte_strings(emu, prefetch):
    w = EO.random_weights(71)
    g = np.random.default_rng(72)
    seqs = [b"", b"A", b"<ACGNNT>", b"ACGTNACGT" * 20]
    for ln in (2, 60, 122, 123, 124, 150, 152, 300):
        s = bytes(g.choice(np.frombuffer(b"ACGTacgt", np.uint8), ln))
        seqs += [s, b"<" + s + b">"]
    assert len(seqs) % 16 != 0 and len(seqs) > 16  # two CTAs, the second partly filled
    got = _embed(emu, w, seqs=seqs)
    assert float(np.abs(got - EO.embed(seqs, w)).max()) < TOL
    # more sequences than one pass of the (emulated 2-SM) device: 2 * 16 * 4 = 128 per pass
    many = [bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), 30)) for _ in range(150)]
    got = _embed(emu, w, max_len=16, seqs=many)
    assert float(np.abs(got - EO.embed(many, w, max_len=16)).max()) < TOL


def test_embed_kernels_window_sources(emu, prefetch):
    w = EO.random_weights(73)
    g = np.random.default_rng(74)
    genome = bytes(g.choice(np.frombuffer(b"ACGTN", np.uint8), 400, p=[0.245, 0.245, 0.245, 0.245, 0.02]))
    cls = _classes(genome)
    ref_len = 150
    # id progression, decorated (what drm_embed_windows does), with ids past the end
    ids = [3 + 29 * i for i in range(19)]
    want = EO.embed([(b"<" + _window(genome, i, ref_len) + b">") if _window(genome, i, ref_len) is not None else b"" for i in ids], w)
    got = _embed(emu, w, cls=cls, form=0, ref_len=ref_len, first_id=3, id_step=29, n=len(ids))
    assert float(np.abs(got - want).max()) < TOL
    # id list, bare windows (the dynamic L2 rerank's source)
    ids = g.integers(0, 2 * (len(genome) - ref_len + 1) + 6, 21)
    want = EO.embed([_window(genome, int(i), ref_len) or b"" for i in ids], w)
    got = _embed(emu, w, cls=cls, form=0, ref_len=ref_len, ids=ids, decorate=0)
    assert float(np.abs(got - want).max()) < TOL
    # id list into a window table, decorated (the static L2 rerank's source)
    rows = [genome[p:p + 40] for p in range(0, 360, 40)]
    table = _classes(b"".join(rows))
    ids = np.array([8, 0, 3, 3, 9, 5], np.uint64)  # 9 = one past the table
    want = EO.embed([(b"<" + rows[int(i)] + b">") if i < len(rows) else b"" for i in ids], w)
    got = _embed(emu, w, cls=table, form=1, ref_len=40, ids=ids, decorate=1)
    assert float(np.abs(got - want).max()) < TOL


def test_l2_rerank_kernel(emu):
    g = np.random.default_rng(75)
    n, U, k = 11, 60, 7
    emb = g.standard_normal((U, 128)).astype(np.float32)
    emb[10:20] = emb[30:40]  # identical embeddings at different rows: equal distances, the order decides
    q = g.standard_normal((n, 128)).astype(np.float32)
    sizes = [0, 3, 7, 8, 40, 33, 64, 7, 12, 100, 20]  # empty, too few, exactly k, more
    starts = np.zeros(n + 1, np.int64); starts[1:] = np.cumsum(sizes)
    mp = g.integers(0, U, int(starts[-1])).astype(np.uint32)
    out_d = np.zeros((n, k), np.float32); out_r = np.zeros((n, k), np.uint32); cnt = np.full(n, 99, np.int64)
    vp = lambda x: x.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = emu.emu_l2_rerank(vp(emb), vp(mp), vp(starts), vp(q), C.c_int64(n), k, C.c_int64(int(starts[-1])), vp(out_d), vp(out_r), vp(cnt))
    assert rc == 0
    for qi in range(n):
        a, b = int(starts[qi]), int(starts[qi + 1])
        if b - a == 0:
            assert cnt[qi] == 0
        elif b - a < k:
            assert cnt[qi] == -(b - a) - 1
        else:
            d = np.array([L2.l2_dist(emb[m], q[qi]) for m in mp[a:b]], np.float32)
            idx = L2.partial_sort_asc(d, k)
            assert cnt[qi] == k
            assert np.array_equal(out_d[qi].view(np.uint32), d[idx].view(np.uint32)), qi
            assert np.array_equal(out_r[qi], mp[a:b][idx]), qi


@full_only
def test_same_checks_under_address_sanitizer():
    """The three tests above once more in a child interpreter with the ASan build preloaded: any out-of-bounds access of
    the kernel code (global or 'shared' arrays) aborts the child."""
    sys.path.insert(0, os.path.join(ROOT, "tools", "cpu_emu"))
    import emu_build
    lib = emu_build.build(sanitize=True)
    asan = subprocess.run(["/usr/bin/g++", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
    if not os.path.isabs(asan) or not os.path.exists(asan):
        pytest.skip("libasan not found")
    code = (
        "import sys, ctypes as C; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import test_emu_kernels as T\n"
        "lib = C.CDLL(%r); lib.emu_last_error.restype = C.c_char_p\n"
        "import os\n"
        "os.environ['DRM_EMBED_PREFETCH'] = '0'; T.test_embed_kernels_byte_strings(lib, '0')\n"
        "os.environ['DRM_EMBED_PREFETCH'] = '1'; T.test_embed_kernels_window_sources(lib, '1')\n"
        "T.test_l2_rerank_kernel(lib)\n"
        "print('asan run ok')\n" % (ROOT, os.path.join(ROOT, "tests"), lib))
    env = dict(os.environ, LD_PRELOAD=asan, ASAN_OPTIONS="detect_leaks=0:abort_on_error=0")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0 and "asan run ok" in r.stdout, (r.stdout[-1500:], r.stderr[-3000:])


@full_only
def test_l2_stage_through_the_c_abi_on_the_emulation():
    """tests/l2_gpu_check.py - the first-hardware-run check of drm_l2_rerank - executed against the emulated library:
    the real api.cu (compiled as C++), l2_plan.cpp, and the embedder / L2 kernel sources on host threads, behind the same
    Python mirror.  Everything after the embeddings must equal the oracle bit for bit."""
    sys.path.insert(0, os.path.join(ROOT, "tools", "cpu_emu"))
    import emu_build
    lib = emu_build.build(sanitize=False)
    code = (
        "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import deepreadmapper_b200.api as A\n"
        "A._LIB_PATH = %r; A._lib = None\n"
        "import l2_gpu_check\n"
        "l2_gpu_check.main(4)\n" % (ROOT, os.path.join(ROOT, "tests"), lib))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0 and "device == oracle" in r.stdout, (r.stdout[-1500:], r.stderr[-3000:])


@full_only
def test_pipeline_l2_stage_host_code_on_the_emulation():
    """tests/l2_pipeline_check.py on the emulated library: the C++ host code behind `pipeline`'s DRM_RERANK=l2
    (post_process_l2_flat via `formats_tool l2flat`, LD_PRELOAD = the emulation) equals the oracle bit for bit, dynamic and
    static (bare lookup-mode windows, device mode 2)."""
    if not os.path.exists(os.path.join(ROOT, "build", "bin", "formats_tool")):
        pytest.skip("host binaries not built (python -c 'import __graft_entry__ as g; g.build()')")
    sys.path.insert(0, os.path.join(ROOT, "tools", "cpu_emu"))
    import emu_build
    lib = emu_build.build(sanitize=False)
    code = (
        "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import deepreadmapper_b200.api as A\n"
        "A._LIB_PATH = %r; A._lib = None\n"
        "import l2_pipeline_check\n"
        "l2_pipeline_check.main(4, %r)\n" % (ROOT, os.path.join(ROOT, "tests"), lib, lib))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0 and "host code == oracle" in r.stdout, (r.stdout[-1500:], r.stderr[-3000:])


def _run_child(code, sanitize):
    sys.path.insert(0, os.path.join(ROOT, "tools", "cpu_emu"))
    import emu_build
    lib = emu_build.build(sanitize=sanitize)
    env = dict(os.environ)
    if sanitize:
        asan = subprocess.run(["/usr/bin/g++", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
        if not os.path.isabs(asan) or not os.path.exists(asan):
            pytest.skip("libasan not found")
        env.update(LD_PRELOAD=asan, ASAN_OPTIONS="detect_leaks=0:abort_on_error=0")
    head = ("import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
            "import deepreadmapper_b200.api as A\n"
            "A._LIB_PATH = %r; A._lib = None\n" % (ROOT, os.path.join(ROOT, "tests"), lib))
    return subprocess.run([sys.executable, "-c", head + code], capture_output=True, text=True, env=env, timeout=1500)


def test_align_merge_and_input_kernels_on_the_emulation():
    """csrc/align_kernel.cu, csrc/merge_kernel.cu (warp shuffles, dynamic shared memory) and csrc/input_kernels.cu through the C
    ABI on host threads: tests/emu_more_checks.py.  DRM_EMU_FULL=1: on the AddressSanitizer build, plus the (slow) input kernels."""
    code = "import emu_more_checks as M\nM.check_align(); M.check_merge()\n" + ("M.check_input()\n" if FULL else "print('input kernels ok')\n")
    r = _run_child(code, sanitize=FULL)
    assert r.returncode == 0 and all(x in r.stdout for x in ("align kernel ok", "merge kernel ok", "input kernels ok")), (r.stdout[-1500:], r.stderr[-3000:])


def test_search_kernels_on_the_emulation():
    """csrc/search_exact.cu (the default, FAISS-exact kernel), csrc/search_kernel.cu (canonical-order kernels) and
    csrc/index_image.cu on host threads: tests/emu_search_check.py.  The exact kernel must equal the oracle's FAISS-rule search
    (tie_rule 0) bit for bit - ids and fp32 distances - the others the canonical rule.  DRM_EMU_FULL=1: every case, under
    AddressSanitizer; default: three cases on the plain build."""
    r = _run_child("import emu_search_check as M\nM.main(%s)\n" % ("" if FULL else "n_reads=5, light=True"), sanitize=FULL)
    assert r.returncode == 0 and "search kernels == oracle" in r.stdout, (r.stdout[-1500:], r.stderr[-3000:])
