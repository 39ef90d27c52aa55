"""Throughput of the embedder with its input resident on the device.

    python tools/embed_bench.py [--reads N] [--reps R] [--out gpurun_out/embed_bench.json]

Random weights of the model's architecture (throughput does not depend on their values), 150-base reads held as
'<' + bases + '>' like drm_reads does, and the same number of reference windows taken by id from a random genome.
Timed with CUDA events on the stream the kernels are queued on (the legacy default stream, which torch's default
stream is).  Algorithmic work: 12.1 M fp32 FMA per sequence (embed_kernel.cu header)."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepreadmapper_b200 as drm  # noqa: E402


def random_weights(seed):
    g = np.random.default_rng(seed)

    def t(*shape):
        return (g.standard_normal(shape) * 0.12).astype(np.float16)

    return {"emb": t(7638, 64), "W1": t(2, 192, 64), "R1": t(2, 192, 64), "B1": t(2, 256), "W2": t(2, 192, 128),
            "R2": t(2, 192, 64), "B2": t(2, 256)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=148 * 64 * 16)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--out", default="gpurun_out/embed_bench.json")
    a = ap.parse_args()
    import torch
    dev = torch.device("cuda:0")
    n = a.reads
    g = np.random.default_rng(3)
    bases = g.choice(np.frombuffer(b"ACGT", np.uint8), (n, 152))
    bases[:, 0] = ord("<"); bases[:, 151] = ord(">")
    flat = torch.from_numpy(bases.reshape(-1)).to(dev)
    off = (torch.arange(n + 1, dtype=torch.int64) * 152).to(dev)
    out = torch.empty((n, 128), dtype=torch.float32, device=dev)
    v = drm.Vectorizer(weights=random_weights(1))
    genome = bytes(g.choice(np.frombuffer(b"ACGT", np.uint8), n // 2 + 149))
    ref = drm.Reference(genome=genome, ref_len=150)

    def timed(fn):
        fn(); torch.cuda.synchronize()
        ms = []
        for _ in range(a.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        return float(np.median(ms))

    # the fp32 (CUDA-core) kernels: without and with the early load of a step's input projections
    os.environ["DRM_EMBED_TC"] = "0"
    os.environ["DRM_EMBED_PREFETCH"] = "0"
    ms_reads = timed(lambda: v.vectorize_device(flat.data_ptr(), off.data_ptr(), 0, n, out.data_ptr()))
    chk = float(out.abs().mean().item())
    os.environ["DRM_EMBED_PREFETCH"] = "1"
    ms_reads_pf = timed(lambda: v.vectorize_device(flat.data_ptr(), off.data_ptr(), 0, n, out.data_ptr()))
    chk_pf = float(out.abs().mean().item())
    ref_out = out.clone()
    ms_win = timed(lambda: v.vectorize_windows_device(ref, 0, 1, n, out.data_ptr()))
    # the tensor-core form (tcgen05 + TMEM; gru_tc_kernel)
    os.environ["DRM_EMBED_TC"] = "1"
    ms_reads_tc = timed(lambda: v.vectorize_device(flat.data_ptr(), off.data_ptr(), 0, n, out.data_ptr()))
    chk_tc = float(out.abs().mean().item())
    tc_max_diff = float((out - ref_out).abs().max().item())
    ms_win_tc = timed(lambda: v.vectorize_windows_device(ref, 0, 1, n, out.data_ptr()))
    os.environ.pop("DRM_EMBED_TC")
    os.environ.pop("DRM_EMBED_PREFETCH")
    fma = 12.1e6
    res = {"sequences": n, "reads_ms": ms_reads, "reads_per_s": n / ms_reads * 1e3, "windows_ms": ms_win,
           "windows_per_s": n / ms_win * 1e3, "fp32_tflops_reads": 2 * fma * n / ms_reads * 1e3 / 1e12,
           "mean_abs_output": chk, "prefetch_reads_ms": ms_reads_pf, "prefetch_reads_per_s": n / ms_reads_pf * 1e3,
           "prefetch_mean_abs_output": chk_pf,
           "tc_reads_ms": ms_reads_tc, "tc_reads_per_s": n / ms_reads_tc * 1e3, "tc_windows_ms": ms_win_tc,
           "tc_windows_per_s": n / ms_win_tc * 1e3, "tc_mean_abs_output": chk_tc, "tc_max_abs_diff_vs_fp32": tc_max_diff,
           "tc_equivalent_tflops_reads": 2 * fma * n / ms_reads_tc * 1e3 / 1e12,
           "gpu": torch.cuda.get_device_name(0)}
    print(json.dumps(res))
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
