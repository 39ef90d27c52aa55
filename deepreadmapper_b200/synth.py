"""Seeded synthetic inputs for the hot path: genome, reads and SURROGATE embeddings.

The reference's embedder (OpenVINO bi-GRU, src/inference/*) is outside the path and unavailable, so
reference windows and reads are embedded by the same cheap stand-in: 4-mer counts of the first 122
bases (Config::Inference::MAX_LEN = 123 truncation) times a fixed Gaussian projection to 128-d.  It
keeps what drives HNSW-PQ cost: adjacent windows are near-duplicates, both strands are indexed, reads
land next to their true window.  Window ids follow the reference: id = 2 * position + strand
(src/utils/parse_inputs.cpp:351-355).  Host-side tooling only - never on the timed path.
"""
from __future__ import annotations

import numpy as np

DIM = 128
KMER = 4
EMB_BASES = 122
N_KMERS = EMB_BASES - KMER + 1  # 119
_BASES = np.frombuffer(b"ACGT", dtype=np.uint8)


def projection(seed: int = 1234) -> np.ndarray:
    rng = np.random.default_rng(seed)
    return (rng.standard_normal((DIM, 256)) / 16.0).astype(np.float32)


def random_genome(n_bases: int, seed: int = 42) -> bytes:
    rng = np.random.default_rng(seed)
    return _BASES[rng.integers(0, 4, size=n_bases, dtype=np.uint8)].tobytes()


def _codes(seq_u8: np.ndarray) -> np.ndarray:
    """ACGT -> 0..3 (anything else -> 0)."""
    lut = np.zeros(256, dtype=np.uint8)
    lut[ord("C")] = 1
    lut[ord("G")] = 2
    lut[ord("T")] = 3
    return lut[seq_u8]


def _kmer_codes(c: np.ndarray) -> np.ndarray:
    """4-mer code at every position of a 0..3 code array (last axis)."""
    return (c[..., 0:-3].astype(np.int32) << 6) | (c[..., 1:-2].astype(np.int32) << 4) | (c[..., 2:-1].astype(np.int32) << 2) | c[..., 3:].astype(np.int32)


_RC_KMER = np.array([sum((3 - ((k >> (2 * i)) & 3)) << (2 * (3 - i)) for i in range(4)) for k in range(256)], dtype=np.int32)


def embed_windows(genome: bytes, ref_len: int = 150, stride: int = 1, chunk: int = 1 << 18, proj=None) -> np.ndarray:
    """Embeddings of every reference window, interleaved forward / reverse-complement: row 2w (+1) is the
    window starting at base w*stride (its reverse complement).  Sliding sums via a float64 prefix sum."""
    P = projection() if proj is None else proj
    Pt = P.T.astype(np.float64)  # 256 x 128
    g = _codes(np.frombuffer(genome, dtype=np.uint8))
    L = len(g)
    n_pos = (L - ref_len) // stride + 1 if L >= ref_len else 0
    out = np.empty((2 * n_pos, DIM), dtype=np.float32)
    K = _kmer_codes(g)  # L - 3
    for w0 in range(0, n_pos, chunk):
        w1 = min(n_pos, w0 + chunk)
        b0, b1 = w0 * stride, (w1 - 1) * stride + ref_len  # bases covered
        kk = K[b0:b1 - 3]
        cf = np.zeros((len(kk) + 1, DIM), dtype=np.float64)
        np.cumsum(Pt[kk], axis=0, out=cf[1:])
        cr = np.zeros((len(kk) + 1, DIM), dtype=np.float64)
        np.cumsum(Pt[_RC_KMER[kk]], axis=0, out=cr[1:])
        s = (np.arange(w0, w1) * stride) - b0
        out[2 * w0:2 * w1:2] = (cf[s + N_KMERS] - cf[s]).astype(np.float32)
        # reverse complement: its first 122 bases are the rc of the window's last 122 bases
        t = s + (ref_len - EMB_BASES)
        out[2 * w0 + 1:2 * w1:2] = (cr[t + N_KMERS] - cr[t]).astype(np.float32)
    return out


def embed_windows_torch(genome: bytes, device, ref_len: int = 150, stride: int = 1, chunk: int = 1 << 20, return_device: bool = False):
    """Same as embed_windows with the prefix sums on a torch device (data generation for bench.py; equal to the
    numpy version up to float32 rounding of the float64 sums).  return_device=True leaves the result on the device
    (a torch tensor) for the GPU index builder."""
    import torch
    Pt = torch.from_numpy(projection().T.astype(np.float64)).to(device)
    g = _codes(np.frombuffer(genome, dtype=np.uint8))
    L = len(g)
    n_pos = (L - ref_len) // stride + 1 if L >= ref_len else 0
    out = torch.empty((2 * n_pos, DIM), dtype=torch.float32, device=device) if return_device else np.empty((2 * n_pos, DIM), dtype=np.float32)
    K = torch.from_numpy(_kmer_codes(g).astype(np.int64)).to(device)
    rck = torch.from_numpy(_RC_KMER.astype(np.int64)).to(device)
    for w0 in range(0, n_pos, chunk):
        w1 = min(n_pos, w0 + chunk)
        b0, b1 = w0 * stride, (w1 - 1) * stride + ref_len
        kk = K[b0:b1 - 3]
        zero = torch.zeros((1, DIM), dtype=torch.float64, device=device)
        cf = torch.cat([zero, torch.cumsum(Pt[kk], dim=0)])
        cr = torch.cat([zero, torch.cumsum(Pt[rck[kk]], dim=0)])
        s = torch.arange(w0, w1, device=device) * stride - b0
        t = s + (ref_len - EMB_BASES)
        blk = torch.empty((2 * (w1 - w0), DIM), dtype=torch.float32, device=device)
        blk[0::2] = (cf[s + N_KMERS] - cf[s]).float()
        blk[1::2] = (cr[t + N_KMERS] - cr[t]).float()
        out[2 * w0:2 * w1] = blk if return_device else blk.cpu().numpy()
        del cf, cr, blk
    return out


def embed_reads(reads: np.ndarray, chunk: int = 1 << 16, proj=None) -> np.ndarray:
    """reads: (n, L) uint8 ASCII, L >= 122."""
    P = projection() if proj is None else proj
    n = reads.shape[0]
    out = np.empty((n, DIM), dtype=np.float32)
    for i0 in range(0, n, chunk):
        c = _codes(reads[i0:i0 + chunk, :EMB_BASES])
        kk = _kmer_codes(c)  # m x 119
        m = kk.shape[0]
        flat = (np.arange(m, dtype=np.int64)[:, None] * 256 + kk).reshape(-1)
        cnt = np.bincount(flat, minlength=m * 256).reshape(m, 256).astype(np.float32)
        out[i0:i0 + m] = cnt @ P.T
    return out


_COMP = np.zeros(256, dtype=np.uint8)
for _a, _b in zip(b"ACGTN", b"TGCAN"):
    _COMP[_a] = _b


def reverse_complement(seq: bytes) -> bytes:
    return _COMP[np.frombuffer(seq, dtype=np.uint8)][::-1].tobytes()


def simulate_reads(genome: bytes, n: int, read_len: int = 150, sub_rate: float = 0.03, indel_rate: float = 0.002, seed: int = 7):
    """(reads uint8 (n, read_len), true window ids int64).  Uniform position and strand, substitutions at
    sub_rate per base, at most one single-base indel per read (probability read_len * indel_rate)."""
    rng = np.random.default_rng(seed)
    g = np.frombuffer(genome, dtype=np.uint8)
    L = len(g)
    pos = rng.integers(0, L - read_len - 1, size=n)
    strand = rng.integers(0, 2, size=n)
    idx = pos[:, None] + np.arange(read_len + 1)[None, :]
    src = g[idx]  # n x (read_len + 1)
    has = rng.random(n) < read_len * indel_rate
    ipos = rng.integers(10, read_len - 10, size=n)
    is_del = rng.random(n) < 0.5
    reads = src[:, :read_len].copy()
    rows = np.nonzero(has & is_del)[0]
    for r in rows:  # deletion: drop base ipos, pull the next genome base in
        p = ipos[r]
        reads[r, p:] = src[r, p + 1:read_len + 1]
    rows = np.nonzero(has & ~is_del)[0]
    ins = _BASES[rng.integers(0, 4, size=n)]
    for r in rows:  # insertion: new base at ipos, last base falls off
        p = ipos[r]
        reads[r, p + 1:] = src[r, p:read_len - 1]
        reads[r, p] = ins[r]
    sub = rng.random((n, read_len)) < sub_rate
    shift = rng.integers(1, 4, size=(n, read_len)).astype(np.uint8)
    codes = _codes(reads)
    reads = np.where(sub, _BASES[(codes + shift) & 3], reads)
    rc = strand == 1
    reads[rc] = _COMP[reads[rc]][:, ::-1]
    ids = 2 * pos.astype(np.int64) + strand
    return np.ascontiguousarray(reads), ids
