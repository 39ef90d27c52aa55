#!/usr/bin/env python
"""One config-1-sized pass over every kernel family, for `compute-sanitizer --tool memcheck|racecheck|synccheck`
(SURVEY.md §5): search (register pool with / without the visited cache, shared-memory pool, pair), SW rerank (profile,
PRMT, generic) + selection + expansion, the GPU index builder (insert + back-link kernels), the per-read K-way merge,
the SW traceback and the FASTA / FASTQ kernels.  Results are compared with the oracle so that a sanitizer-clean run is
also a correct one.  usage: compute-sanitizer --tool racecheck python tools/sanitize_case.py [--part NAME]"""
import argparse
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--part", default="all", help="all | search | rerank | build | merge | align | input | l2")
    ap.add_argument("--genome", type=int, default=1000)
    ap.add_argument("--reads", type=int, default=48)
    a = ap.parse_args()
    import deepreadmapper_b200 as drm
    from deepreadmapper_b200 import synth
    from oracle import bindings as ob

    want = lambda p: a.part in ("all", p)
    g = synth.random_genome(a.genome, 42)
    reads, truth = synth.simulate_reads(g, a.reads)
    Q = synth.embed_reads(reads)
    seqs = [bytes(r) for r in reads]
    td = tempfile.mkdtemp()
    path = os.path.join(td, "c1.index")
    E = synth.embed_windows(g)

    if want("build"):
        drm.build_index_gpu(E, path, efc=64, seed=3)
        oix = ob.OracleIndex(path)
        D0, I0, _ = oix.search(Q, 16, 32, tie_rule=1)
        assert (I0[:, 0] >= 0).all()
        print("build: gpu builder ok, n =", E.shape[0], flush=True)
    else:
        drm.build_index(E, path, efc=64, seed=3)

    if want("search"):
        oix = ob.OracleIndex(path)
        for kern, vis, (k, ef) in [("exact", "512", (128, 128)), ("exact", "0", (32, 64)), ("exact", "512", (200, 128)), ("reg", "512", (128, 128)), ("reg", "0", (32, 64)), ("smem", "512", (200, 128)), ("pair", "512", (64, 64)),
                                   ("reg", "512", (8, 200))]:
            os.environ["DRM_SEARCH_KERNEL"], os.environ["DRM_SEARCH_VIS"] = kern, vis
            ix = drm.Index(path, devices=[0])
            D, I = ix.search(Q, k, ef)
            ix.close()
            D0, I0, _ = oix.search(Q, k, ef, tie_rule=0 if kern == "exact" else 1)
            assert np.array_equal(I, I0) and np.array_equal(D.view(np.uint32), D0.view(np.uint32)), (kern, vis, k, ef)
            print(f"search: {kern} vis={vis} k={k} ef={ef} ok", flush=True)
        os.environ.pop("DRM_SEARCH_KERNEL"), os.environ.pop("DRM_SEARCH_VIS")

    if want("rerank"):
        oix = ob.OracleIndex(path)
        _, I0, _ = oix.search(Q, 32, 64, tie_rule=0)
        seqs2 = list(seqs)
        seqs2[1] = seqs2[1][:70] + b"NNN" + seqs2[1][73:]
        seqs2[2] = b"<" + seqs2[2] + b">"
        import bench
        wins = bench.window_table(g)
        for variant in ("default", "prmt", "blend2"):
            if variant == "default":
                os.environ.pop("DRM_SW_KERNEL", None)
            else:
                os.environ["DRM_SW_KERNEL"] = variant
            ref = drm.Reference(genome=g, ref_len=150, devices=[0])
            for mode, stride, k, kc in [(1, 1, 32, 32), (0, 1, 10, 32), (1, 4, 40, 16), (0, 8, 64, 8)]:
                nb = I0[:, :kc] // stride if stride > 1 else I0[:, :kc]
                refb, nref = (wins, len(wins) // 150) if mode == 0 else (g, len(g))
                rc, s0, i0, _ = ob.post_process_sw(mode, nb, refb, nref, 150, seqs2, stride, k, kc)
                try:
                    s, ids = ref.rerank(mode, nb, seqs2, stride, k, kc)
                except drm.DrmError:
                    assert rc != 0, (variant, mode, stride)
                    continue
                assert rc == 0 and np.array_equal(s, s0) and np.array_equal(ids, i0), (variant, mode, stride)
            ref.close()
            print(f"rerank: {variant} ok", flush=True)
        os.environ.pop("DRM_SW_KERNEL", None)
        a_, b_ = [s[:100] for s in seqs], [s[20:] + b"x" for s in seqs]
        assert drm.sw_pairs(a_, b_).tolist() == ob.sw_pairs(a_, b_).tolist()
        print("rerank: generic pairs ok", flush=True)

    if want("merge"):
        import torch
        rng = np.random.default_rng(1)
        S, n, k = 4, 40, 32
        D = np.sort(rng.integers(0, 20, (S, n, k)).astype(np.float32), axis=2)
        I = rng.permutation(S * n * k).reshape(S, n, k).astype(np.int64)
        for s in range(S):
            for r in range(n):
                o = np.lexsort((I[s, r].astype(np.uint64), D[s, r]))
                D[s, r], I[s, r] = D[s, r][o], I[s, r][o]
        dD, dI = torch.from_numpy(D).cuda(), torch.from_numpy(I).cuda()
        oD, oI = torch.empty((n, k), dtype=torch.float32, device="cuda"), torch.empty((n, k), dtype=torch.int64, device="cuda")
        drm.merge_topk_device(0, dD.data_ptr(), dI.data_ptr(), S, n, k, oD.data_ptr(), oI.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        for r in range(n):
            d, i = D[:, r].reshape(-1), I[:, r].reshape(-1)
            o = np.lexsort((i.astype(np.uint64), d))[:k]
            assert np.array_equal(oD[r].cpu().numpy(), d[o]) and np.array_equal(oI[r].cpu().numpy(), i[o])
        print("merge: ok", flush=True)

    if want("align"):
        ref = drm.Reference(genome=g, ref_len=150, devices=[0])
        ids = np.stack([truth, truth ^ 1], axis=1).astype(np.uint64)
        out = ref.align(ids, seqs)
        ref.close()
        print("align: ok,", len(out["score"]) if isinstance(out, dict) else "done", flush=True)

    if want("input"):
        fa = b"> hdr\n" + b"\n".join(g[i:i + 70] for i in range(0, len(g), 70)) + b"\n"
        assert drm.fasta_clean(fa) == g
        fq = b"".join(b"@r%d\n%s\n+\n%s\n" % (i, s, b"I" * len(s)) for i, s in enumerate(seqs))
        r = drm.Reads(fq, devices=[0])
        assert [x for x in r.sequences()] == [b"<" + s + b">" for s in seqs]
        r.close()
        print("input: ok", flush=True)
    print("sanitize_case: all requested parts ok", flush=True)


if __name__ == "__main__":
    main()
