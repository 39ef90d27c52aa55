#!/usr/bin/env python
"""GPU index builder vs host builder on the bench workload (BASELINE.json configs[1] index: 9.28 M vectors, EFC 200).

Builds the index with drm_build_index_gpu_device (stage times), optionally with the host builder too (--cpu, ~140 s on
16 cores), and for each file reports: graph statistics, the search kernel's time and counters on --reads reads at
EF = K = 128, recall@K against the exact ADC scan on --sample reads, and the share of reads whose planted window is
among the K labels.  Measurement tool: the oracle is used as the checker only."""
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
    from deepreadmapper_b200 import synth
    from oracle import bindings as ob

    ap = argparse.ArgumentParser()
    ap.add_argument("--genome-bases", type=int, default=bench.GENOME_BASES)
    ap.add_argument("--efc", type=int, default=200)
    ap.add_argument("--reads", type=int, default=262144)
    ap.add_argument("--sample", type=int, default=2000)
    ap.add_argument("--cpu", action="store_true", help="also build with the host builder (or reuse the bench's cached file)")
    ap.add_argument("--cache", default=os.environ.get("DRM_CACHE", os.path.join("/tmp", "drm_b200_cache")))
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    os.makedirs(a.cache, exist_ok=True)
    torch.cuda.set_device(0)
    genome = bench.make_genome(a)
    t0 = time.time()
    E = synth.embed_windows_torch(genome, "cuda:0", return_device=True)
    torch.cuda.synchronize()
    bench.log(f"embedded {E.shape[0]} windows in {time.time() - t0:.1f}s")
    files = {}
    gpath = os.path.join(a.cache, f"gpu_built_g{a.genome_bases}_efc{a.efc}.index")
    t0 = time.time()
    tim = drm.build_index_gpu_device(E.data_ptr(), E.shape[0], E.shape[1], gpath, train_step=2, efc=a.efc, seed=1234)
    files["gpu"] = dict(path=gpath, build_s=time.time() - t0, stages_s=tim)
    bench.log(f"GPU build: {files['gpu']['build_s']:.1f}s wall incl. file write; stages {tim}")
    if a.cpu:
        t0 = time.time()
        a.builder = "cpu"
        cpath, _ = bench.ensure_index(a, genome, "cuda:0")
        files["cpu"] = dict(path=cpath, build_s=time.time() - t0)
    del E
    torch.cuda.empty_cache()

    n, ns = a.reads, min(a.sample, a.reads)
    reads, truth, Q = bench.make_reads(genome, n, seed=7)
    dQ = torch.from_numpy(Q).cuda()
    K, EF = bench.K, bench.EF
    dD = torch.empty((n, K), dtype=torch.float32, device="cuda")
    dI = torch.empty((n, K), dtype=torch.int64, device="cuda")
    dcnt = torch.zeros(3, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream()
    out = {}
    for name, f in files.items():
        oix = ob.OracleIndex(f["path"])
        assert oix.validate() == 0, (name, oix.validate())
        _, _, deg = oix.arrays()
        ix = drm.Index(f["path"], devices=[0])
        ts = []
        for it in range(5):
            dcnt.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            ix.search_device(dQ.data_ptr(), n, K, EF, dD.data_ptr(), dI.data_ptr(), dcnt.data_ptr(), stream.cuda_stream)
            e1.record(stream)
            torch.cuda.synchronize()
            if it >= 2:
                ts.append(e0.elapsed_time(e1))
        cnt = (dcnt.cpu().numpy() / n).tolist()
        I = dI[:ns].cpu().numpy()
        _, Ie = oix.adc_scan(Q[:ns], K, threads=bench.host_threads())
        rec = float(np.mean([len(set(I[i].tolist()) & set(Ie[i].tolist())) / K for i in range(ns)]))
        Iall = dI.cpu().numpy()
        planted = float(np.mean(((Iall >> 1) == (truth[:, None] >> 1)).any(axis=1)))
        near = float(np.mean((np.abs((Iall >> 1) - (truth[:, None] >> 1)) <= 2).any(axis=1)))
        ix.close()
        out[name] = dict(build_s=f["build_s"], stages_s=f.get("stages_s"), graph=deg, max_level=oix.max_level, search_ms=float(np.mean(ts)),
                         reads=n, nhops0=cnt[0], nhops_up=cnt[1], ndis_eval=cnt[2], recall_at_k_vs_exact_adc=rec,
                         planted_window_in_top_k=planted, planted_window_within_2bp_in_top_k=near)
        bench.log(f"{name}: {out[name]}")
    print(json.dumps(out), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
