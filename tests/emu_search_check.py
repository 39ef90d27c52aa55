"""The search kernels' SOURCE on host threads (tools/cpu_emu) against the oracle - run in a child interpreter by
tests/test_emu_kernels.py with deepreadmapper_b200.api pointed at the emulated library.

  * search_exact.cu (default kernel) == oracle tie_rule 0 (FAISS's own tie behaviour), ids and fp32 distance bits, for
    k == ef, k < ef, k > ef (count_below stop rule), k not a multiple of 32, with and without visited cache / read-ahead;
  * search_kernel.cu (canonical-order kernels: register pool, shared-memory pool, pair) == oracle tie_rule 1;
  * index_image.cu builds the fused-row image both use.
"""
import os
import tempfile

import numpy as np


def main(n_reads=16, genome_bases=1000, light=False):
    import deepreadmapper_b200 as drm
    from deepreadmapper_b200 import synth
    from oracle import bindings as ob

    g = synth.random_genome(genome_bases, 42)
    path = os.path.join(tempfile.mkdtemp(), "emu.index")
    drm.build_index(synth.embed_windows(g), path, efc=64, seed=3)
    reads, _ = synth.simulate_reads(g, n_reads)
    Q = synth.embed_reads(reads)
    oix = ob.OracleIndex(path)
    ix = drm.Index(path, devices=[0])
    cases = [("exact", "512", "1", (128, 128)), ("exact", "0", "0", (8, 32)), ("exact", "512", "1", (200, 128)), ("exact", "1024", "1", (100, 128)),
             ("exact", "512", "1", (64, 300)), ("reg", "512", "1", (128, 128)), ("smem", "512", "1", (200, 128)), ("pair", "512", "1", (32, 64))]
    if light:
        cases = [cases[0], cases[2]]  # exact kernel: k == ef and k > ef (count_below); the canonical kernels run under DRM_EMU_FULL=1
    for kern, vis, ra, (k, ef) in cases:
        os.environ.update(DRM_SEARCH_KERNEL=kern, DRM_SEARCH_VIS=vis, DRM_SEARCH_RA=ra)
        D1, I1, st1 = ix.search(Q, k, ef, return_stats=True)
        D0, I0, st0 = oix.search(Q, k, ef, tie_rule=0 if kern == "exact" else 1)
        assert np.array_equal(I0, I1) and np.array_equal(D0.view(np.uint32), D1.view(np.uint32)), (kern, vis, ra, k, ef)
        assert st1["nhops0"] == st0["nhops"] and st1["nhops_up"] == st0["nhops_up"] and st1["ndis_eval"] >= st0["ndis"]
    ix.close()
    for v in ("DRM_SEARCH_KERNEL", "DRM_SEARCH_VIS", "DRM_SEARCH_RA"):
        os.environ.pop(v, None)
    print("search kernels == oracle")


if __name__ == "__main__":
    main()
