"""Run by tests/test_emu_kernels.py in a child interpreter whose Python mirror points at the emulated library
(tools/cpu_emu): the traceback kernel (csrc/align_kernel.cu), the shard-merge kernel (csrc/merge_kernel.cu) and the FASTA /
FASTQ kernels (csrc/input_kernels.cu) - warp shuffles, atomics, dynamic shared memory - on host threads, against the oracle / numpy.  Both kernels pass on hardware
(tests/test_gpu_align.py, tests/test_gpu_parity.py::test_merge_topk); here their source additionally runs under
AddressSanitizer when the sanitised library is preloaded."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import deepreadmapper_b200 as drm  # noqa: E402
from oracle import bindings as ob  # noqa: E402

COMP = bytes.maketrans(b"ACGTN", b"TGCAN")


def window(g, wid, ref_len):
    pos, rc = wid >> 1, wid & 1
    if pos + ref_len > len(g):
        return b""
    w = g[pos:pos + ref_len]
    return w.translate(COMP)[::-1] if rc else w


def check_align():
    rng = np.random.default_rng(81)
    g = bytes(rng.choice(np.frombuffer(b"ACGTN", np.uint8), 700, p=[0.245, 0.245, 0.245, 0.245, 0.02]))
    ref_len, max_ops = 150, 160
    ref = drm.Reference(genome=g, ref_len=ref_len)
    nwin = 2 * (len(g) - ref_len + 1)
    reads, ids = [], []
    for i in range(14):
        p = int(rng.integers(0, len(g) - ref_len))
        r = bytearray(g[p:p + ref_len])
        for j in rng.integers(0, ref_len, 6):
            r[j] = b"ACGT"[int(rng.integers(0, 4))]
        if i % 3 == 0:
            del r[40:43]
        if i % 4 == 0:
            r[90:90] = b"GT"
        reads.append(bytes(r))
        ids.append([2 * p, 2 * p + 1, int(rng.integers(0, nwin + 4))])  # the planted window, its reverse strand, a random id (maybe past the end)
    reads += [b"A", b"", b"NNNNACGT", (g + g)[:256]]
    ids += [[0, 1, 2]] * 4
    ids = np.array(ids, np.uint64)
    for strip, seqs in ((0, reads), (1, [b"<" + s + b">" for s in reads])):
        got = ref.align(ids, seqs, strip=strip, max_ops=max_ops)
        for i in range(len(reads)):
            for s in range(ids.shape[1]):
                want = ob.sw_align(window(g, int(ids[i, s]), ref_len), reads[i], max_ops)
                for k in ("score", "ref_beg", "ref_end", "q_beg", "q_end", "n_ops", "nm"):
                    assert int(got[k][i, s]) == want[k], (strip, i, s, k, int(got[k][i, s]), want)
                if want["n_ops"] <= max_ops:
                    ops = [(int(w) >> 4, "MID"[int(w) & 15]) for w in got["ops"][i, s, :want["n_ops"]]]
                    assert ops == want["ops"], (strip, i, s)
    ref.close()
    print("align kernel ok")


def check_merge():
    rng = np.random.default_rng(82)
    S, n, k = 5, 9, 32
    D = np.sort(rng.integers(0, 20, (S, n, k)).astype(np.float32), axis=2)
    I = rng.permutation(S * n * k).reshape(S, n, k).astype(np.int64)
    D[3, 5, 20:] = np.finfo(np.float32).max
    I[3, 5, 20:] = -1
    for s in range(S):
        for r in range(n):
            o = np.lexsort((I[s, r].astype(np.uint64), D[s, r]))
            D[s, r], I[s, r] = D[s, r][o], I[s, r][o]
    D = np.ascontiguousarray(D); I = np.ascontiguousarray(I)
    oD = np.zeros((n, k), np.float32); oI = np.zeros((n, k), np.int64)
    drm.merge_topk_device(0, D.ctypes.data, I.ctypes.data, S, n, k, oD.ctypes.data, oI.ctypes.data, 0)
    for r in range(n):
        d, i = D[:, r].reshape(-1), I[:, r].reshape(-1)
        o = np.lexsort((i.astype(np.uint64), d))[:k]
        assert np.array_equal(oD[r], d[o]) and np.array_equal(oI[r], i[o]), r
    print("merge kernel ok")


def check_input():
    """csrc/input_kernels.cu (tile counts, hierarchical scan, ordered writes; shuffles, atomics) on two small files."""
    fasta = b">h with ACGT letters\r\nacgtnRYKM\r\nACGT\r\n>contig2 GATTACA\nTTTTNN"
    body = fasta[fasta.index(b"\n") + 1:]
    want = bytes(c for c in body.upper() if c in b"ACGTN")  # extract_FASTA_sequence: first line dropped, the rest filtered
    assert drm.fasta_clean(fasta) == want, (drm.fasta_clean(fasta), want)
    fastq = b"@r0 x\nACGTN\n+\nIIIII\n@r1\nGG\n+\nII\n@r2\n\n+\n\n@r3\nTTTTTTTT\n+\nIIIIIIII"
    lines = fastq.split(b"\n")
    reads = drm.Reads(fastq)
    assert reads.sequences() == [b"<" + lines[4 * r + 1] + b">" for r in range(4)], reads.sequences()
    reads.close()
    print("input kernels ok")


if __name__ == "__main__":
    check_align()
    check_merge()
    check_input()
