"""Debug helper (GPU): replays tests/test_gpu_parity.py::test_rerank_random_vs_oracle step by step and prints every
row where the GPU rerank and the oracle disagree, together with both versions and which one the reference confirms."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepreadmapper_b200 as drm
from oracle import bindings as ob

g = json.load(open("tests/golden/sw_golden.json"))["fixtures"]
genome = g["genome"]
comp = {"A": "T", "T": "A", "C": "G", "G": "C", "N": "N"}
wins = []
for p in range(len(genome) - 149):
    w = genome[p:p + 150]; wins.append(w); wins.append("".join(comp[c] for c in reversed(w)))
reads = list(g["reads"][:40])
reads[3] = reads[3][1:-1]; reads[5] = reads[5][:97]; reads[7] = reads[7][:60] + "N" * 5 + reads[7][65:]; reads[9] = reads[9].lower()
print("reads with N:", [i for i, r in enumerate(reads) if "N" in r])
rng = np.random.default_rng(3)
refs = {0: drm.Reference(windows=wins, ref_len=150), 1: drm.Reference(genome=genome, ref_len=150)}
gb = genome.encode(); wb = "".join(wins).encode()
for mode in (0, 1):
    for stride, kc, k in [(1, 128, 128), (1, 64, 10), (4, 20, 50), (8, 32, 128), (8, 5, 20), (16, 8, 60), (2, 100, 33)]:
        hi = len(wins) // stride if mode == 0 else len(genome) // stride
        hi = max(hi - (1 if mode == 1 else 0), kc + 1)
        nb = np.stack([rng.permutation(hi)[:kc] for _ in reads]).astype(np.int64)
        refb, nref = (wb, len(wins)) if mode == 0 else (gb, len(gb))
        rc, s1, i1, nc1 = ob.post_process_sw(mode, nb, refb, nref, 150, reads, stride, k, kc)
        if rc != 0:
            continue
        rcr, sr, ir, ncr = ob.post_process_sw(mode, nb, refb, nref, 150, reads, stride, k, kc, use_ref=True)
        for rep in range(3):
            s2, i2, nc2, _ = refs[mode].rerank(mode, nb, reads, stride, k, kc, return_extra=True)
            bad = [r for r in range(len(reads)) if s2[r].tolist() != s1[r].tolist() or i2[r].tolist() != i1[r].tolist()]
            orc_vs_ref = [r for r in range(len(reads)) if sr[r].tolist() != s1[r].tolist() or ir[r].tolist() != i1[r].tolist()]
            print(mode, stride, kc, k, "rep", rep, "gpu!=oracle rows:", bad, "oracle!=reference rows:", orc_vs_ref)
            for r in bad[:2]:
                d = [(j, int(s1[r][j]), int(s2[r][j]), int(i1[r][j]), int(i2[r][j])) for j in range(k) if s1[r][j] != s2[r][j] or i1[r][j] != i2[r][j]]
                print("   row", r, "len", len(reads[r]), "ncands", int(nc1[r]), int(nc2[r]), "ndiff", len(d), d[:8])
                print("   gpu==reference:", s2[r].tolist() == sr[r].tolist() and i2[r].tolist() == ir[r].tolist())
