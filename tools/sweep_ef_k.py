#!/usr/bin/env python
"""BASELINE.json configs[3]: EF/K sweep on the E. coli-sized dense index (the bench's cached index file).

For EF in {32..512} x K in {8..128} with K <= EF: device-timed search reads/s and search+SW-rerank reads/s on the
GPU for --reads reads, and on a --sample prefix: recall@K against the exact ADC top-K (oracle brute-force scan over
all codes), overlap with the FAISS-rule CPU restatement, and that restatement's own reads/s on the host cores.
Test/measurement infrastructure: it calls the oracle as the checker, never inside the GPU-timed region.
Output: one JSON object per (EF, K) on stdout; `--out file` also writes the list.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch

    import bench
    import deepreadmapper_b200 as drm
    from oracle import bindings as ob

    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=100_000)
    ap.add_argument("--sample", type=int, default=2000)
    ap.add_argument("--efs", default="32,64,128,256,512")
    ap.add_argument("--ks", default="8,16,32,64,128")
    ap.add_argument("--genome-bases", type=int, default=bench.GENOME_BASES)
    ap.add_argument("--efc", type=int, default=200)
    ap.add_argument("--cache", default=os.environ.get("DRM_CACHE", os.path.join("/tmp", "drm_b200_cache")))
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--out", default="")
    a = ap.parse_args()

    genome = bench.make_genome(a)
    path, _ = bench.ensure_index(a, genome, "cuda:0")
    ix = drm.Index(path, devices=[0])
    ref = drm.Reference(genome=genome, ref_len=bench.REF_LEN, devices=[0])
    oix = ob.OracleIndex(path)
    n, ns = a.reads, min(a.sample, a.reads)
    reads, truth, Q = bench.make_reads(genome, n, seed=7)
    threads = os.cpu_count() or 1
    kmax = max(int(k) for k in a.ks.split(","))
    t0 = time.perf_counter()
    De, Ie = oix.adc_scan(Q[:ns], kmax, threads=threads)
    bench.log(f"exact ADC scan of {ns} reads: {time.perf_counter() - t0:.1f}s")

    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream
    dQ = torch.from_numpy(Q).cuda()
    dreads = torch.from_numpy(reads.reshape(-1)).cuda()
    doff = torch.from_numpy(np.arange(n + 1, dtype=np.int64) * bench.REF_LEN).cuda()
    rows = []
    for ef in (int(x) for x in a.efs.split(",")):
        for k in (int(x) for x in a.ks.split(",")):
            if k > ef:
                continue
            dD = torch.empty((n, k), dtype=torch.float32, device="cuda")
            dI = torch.empty((n, k), dtype=torch.int64, device="cuda")
            dS = torch.empty((n, k), dtype=torch.int32, device="cuda")
            dIds = torch.empty((n, k), dtype=torch.int64, device="cuda")
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            ts, tw = [], []
            for rep in range(a.reps + 1):
                ev[0].record(stream)
                ix.search_device(dQ.data_ptr(), n, k, ef, dD.data_ptr(), dI.data_ptr(), 0, sp)
                ev[1].record(stream)
                ref.rerank_device(1, dI.data_ptr(), n, k, dreads.data_ptr(), doff.data_ptr(), bench.REF_LEN, 1, k, k,
                                  dS.data_ptr(), dIds.data_ptr(), 0, 0, 0, sp)
                ev[2].record(stream)
                torch.cuda.synchronize()
                if rep:  # first repetition = warm-up
                    ts.append(ev[0].elapsed_time(ev[1]))
                    tw.append(ev[1].elapsed_time(ev[2]))
            I = dI[:ns].cpu().numpy()
            Dg = dD[:ns].cpu().numpy()
            ids_sw = dIds[:, 0].cpu().numpy()
            t0 = time.perf_counter()
            D0, I0, st = oix.search(Q[:ns], k, ef, tie_rule=0, threads=threads)
            cpu_s = time.perf_counter() - t0
            rec_exact = float(np.mean([len(set(I[i].tolist()) & set(Ie[i, :k].tolist())) / k for i in range(ns)]))
            rec_cpu_exact = float(np.mean([len(set(I0[i].tolist()) & set(Ie[i, :k].tolist())) / k for i in range(ns)]))
            # tie-aware: a returned label counts when its distance is within the exact K-th smallest (many windows share PQ codes)
            rec_tie = float(np.mean([(Dg[i] <= De[i, k - 1]).mean() for i in range(ns)]))
            planted_graph = float(np.mean([truth[i] in I[i] for i in range(ns)]))
            planted_exact = float(np.mean([truth[i] in Ie[i, :k] for i in range(ns)]))
            overlap = float(np.mean([len(set(I[i].tolist()) & set(I0[i].tolist())) / k for i in range(ns)]))
            row = {"EF": ef, "K": k, "reads": n,
                   "gpu_search_reads_per_s": n / (np.mean(ts) / 1e3), "gpu_search_sw_reads_per_s": n / ((np.mean(ts) + np.mean(tw)) / 1e3),
                   "gpu_search_ms": float(np.mean(ts)), "gpu_sw_ms": float(np.mean(tw)),
                   "recall_at_k_vs_exact_adc": rec_exact, "recall_at_k_vs_exact_adc_tie_aware": rec_tie,
                   "planted_window_in_topk_graph": planted_graph, "planted_window_in_topk_exact_adc": planted_exact, "cpu_recall_at_k_vs_exact_adc": rec_cpu_exact,
                   "overlap_vs_cpu_faiss_rule": overlap, "sw_top1_is_planted_position": float((ids_sw >> 1 == (truth >> 1)).mean()),
                   "cpu_search_reads_per_s": ns / cpu_s, "cpu_threads": threads, "cpu_sample": ns,
                   "cpu_ndis_per_read": st["ndis"] / ns, "cpu_nhops_per_read": st["nhops"] / ns}
            rows.append(row)
            print(json.dumps(row), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
