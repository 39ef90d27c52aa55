#!/usr/bin/env python
"""Does running the search of chunk c + 1 CONCURRENTLY with the SW rerank of chunk c (two streams) beat running the two
kernels back to back?  The search kernel is latency / issue bound (IPC 0.65), the SW kernel pipe bound (IPC 0.76): their
stall reasons differ, so co-residency on an SM may fill issue slots neither fills alone.  Device-resident inputs, CUDA events.
usage: python tools/overlap_probe.py [--reads N] [--chunks C ...]"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch

    import bench
    import deepreadmapper_b200 as drm

    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=1_000_000)
    ap.add_argument("--chunks", type=int, nargs="+", default=[1, 4, 8, 16])
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--caps", type=int, nargs="*", default=[1, 2, 3])
    ap.add_argument("--out", default="gpurun_out/overlap_probe.json")
    a = ap.parse_args()
    args = argparse.Namespace(genome_bases=bench.GENOME_BASES, efc=200, builder="gpu", cache=os.environ.get("DRM_CACHE", "/tmp/drm_b200_cache"))
    torch.cuda.set_device(0)
    genome = bench.make_genome(args)
    path, _ = bench.ensure_index(args, genome, "cuda:0")
    ix = drm.Index(path, devices=[0])
    ref = drm.Reference(genome=genome, ref_len=150, devices=[0])
    n, K, EF = a.reads, 128, 128
    reads, truth, Q = bench.make_reads(genome, n, seed=7)
    dQ = torch.from_numpy(Q).cuda()
    dreads = torch.from_numpy(reads.reshape(-1)).cuda()
    doff = torch.from_numpy(np.arange(n + 1, dtype=np.int64) * 150).cuda()
    dD = torch.empty((n, K), dtype=torch.float32, device="cuda")
    dI = torch.empty((n, K), dtype=torch.int64, device="cuda")
    dS = torch.empty((n, K), dtype=torch.int32, device="cuda")
    dIds = torch.empty((n, K), dtype=torch.int64, device="cuda")
    s1, s2 = torch.cuda.Stream(priority=-1), torch.cuda.Stream()  # the search stream has the higher priority: its (capped) grid gets its SM slots first
    res = {}
    ref_out = None
    for C in a.chunks:
        bounds = [(n * c // C, n * (c + 1) // C) for c in range(C)]
        for mode in ("serial", "overlap") + tuple(f"overlap_cap{c}" for c in a.caps if C > 1):
            # overlap_capN: the persistent search kernel takes only N CTAs per SM (5 fit), so the SW CTAs of the previous chunk stay resident beside it
            if mode.startswith("overlap_cap"):
                os.environ["DRM_SEARCH_CTAS_PER_SM"] = mode[len("overlap_cap"):]
            else:
                os.environ.pop("DRM_SEARCH_CTAS_PER_SM", None)

            def run():
                evs = []
                for (c0, c1) in bounds:
                    ssearch = s1
                    srer = s2 if mode != "serial" else s1
                    ix.search_device(dQ[c0:c1].data_ptr(), c1 - c0, K, EF, dD[c0:c1].data_ptr(), dI[c0:c1].data_ptr(), 0, ssearch.cuda_stream)
                    if mode != "serial":
                        e = torch.cuda.Event()
                        e.record(ssearch)
                        srer.wait_event(e)
                    ref.rerank_device(1, dI[c0:c1].data_ptr(), c1 - c0, K, dreads.data_ptr(), doff[c0:].data_ptr(), 150, 1, K, K, dS[c0:c1].data_ptr(),
                                      dIds[c0:c1].data_ptr(), 0, 0, 0, srer.cuda_stream)
            run()
            torch.cuda.synchronize()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record(s1)
            for _ in range(a.reps):
                run()
            if mode != "serial":
                e = torch.cuda.Event()
                e.record(s2)
                s1.wait_event(e)
            t1.record(s1)
            torch.cuda.synchronize()
            ms = t0.elapsed_time(t1) / a.reps
            out = (dI.cpu().numpy().copy(), dS.cpu().numpy().copy(), dIds.cpu().numpy().copy())
            if ref_out is None:
                ref_out = out
            same = all(np.array_equal(x, y) for x, y in zip(ref_out, out))
            res[f"chunks{C}_{mode}"] = {"ms": ms, "reads_per_s": n / (ms / 1e3), "identical": bool(same)}
            print(f"chunks={C:3d} {mode:13s} {ms:8.2f} ms  {n / (ms / 1e3) / 1e6:.3f} M reads/s  identical={same}", flush=True)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(res, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
