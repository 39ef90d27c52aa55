"""Debug helper (GPU): runs the failing rerank case on every SW variant and prints where rows differ from the oracle."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepreadmapper_b200 as drm
from oracle import bindings as ob

g = json.load(open("tests/golden/sw_golden.json"))["fixtures"]
genome = g["genome"]
reads = list(g["reads"][:40])
reads[3] = reads[3][1:-1]; reads[5] = reads[5][:97]; reads[7] = reads[7][:60] + "N" * 5 + reads[7][65:]; reads[9] = reads[9].lower()
rng = np.random.default_rng(3)
gb = genome.encode()
for trial in range(3):
    nb = np.stack([rng.permutation(len(genome) - 1)[:128] for _ in reads]).astype(np.int64)
    rc, s1, i1, nc1 = ob.post_process_sw(1, nb, gb, len(gb), 150, reads, 1, 128, 128)
    for v in ["prmt", "profile", "profile16", "blend2", "blend3", "blend4"]:
        os.environ["DRM_SW_KERNEL"] = v
        ref = drm.Reference(genome=genome, ref_len=150)
        for rep in range(2):
            s2, i2 = ref.rerank(1, nb, reads, 1, 128, 128)
            bad = [r for r in range(len(reads)) if s2[r].tolist() != s1[r].tolist() or i2[r].tolist() != i1[r].tolist()]
            print(trial, v, rep, "bad rows:", bad)
            for r in bad[:2]:
                d = [(j, int(s1[r][j]), int(s2[r][j]), int(i1[r][j]), int(i2[r][j])) for j in range(128) if s1[r][j] != s2[r][j] or i1[r][j] != i2[r][j]]
                print("   row", r, "len", len(reads[r]), "ndiff", len(d), d[:6])
        ref.close()
