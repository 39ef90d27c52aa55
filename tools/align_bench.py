#!/usr/bin/env python
"""Traceback kernel throughput: drm_sw_align on the planted windows of synthetic reads, wall clock
through the host-pointer C ABI (copies included), against the oracle definition on a sample."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import deepreadmapper_b200 as drm
    from deepreadmapper_b200 import synth
    from oracle import bindings as ob
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    g = synth.random_genome(4_641_652, 42)
    reads, truth = synth.simulate_reads(g, n, 150, seed=7)
    seqs = [bytes(r) for r in reads]
    ref = drm.Reference(genome=g, ref_len=150)
    ids = np.stack([truth, truth ^ 1], axis=1).astype(np.uint64)
    out = {}
    for rep in range(3):
        t0 = time.perf_counter()
        got = ref.align(ids, seqs, max_ops=32)
        out["wall_ms"] = (time.perf_counter() - t0) * 1e3
    out["alignments"] = int(ids.size)
    out["alignments_per_s_wall"] = ids.size / (out["wall_ms"] / 1e3)
    out["cells_per_s_wall"] = out["alignments_per_s_wall"] * 150 * 150
    comp = bytes.maketrans(b"ACGTN", b"TGCAN")
    t0 = time.perf_counter()
    ns = 2000
    for i in range(ns):
        pos, rc = int(truth[i]) >> 1, int(truth[i]) & 1
        w = g[pos:pos + 150]
        w = w.translate(comp)[::-1] if rc else w
        want = ob.sw_align(w, seqs[i], 32)
        assert want["score"] == got["score"][i, 0] and want["n_ops"] == got["n_ops"][i, 0] and want["ref_beg"] == got["ref_beg"][i, 0]
    out["oracle_alignments_per_s_1thread"] = ns / (time.perf_counter() - t0)
    out["median_planted_score"] = float(np.median(got["score"][:, 0]))
    out["share_with_indel_ops"] = float(np.mean(got["n_ops"][:, 0] > 1))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
