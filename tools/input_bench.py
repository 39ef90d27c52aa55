#!/usr/bin/env python
"""Input side: device FASTQ / FASTA parsing against the reference's host parsers on the same bytes.
Prints one JSON object.  Measurement tool: oracle/_ref is the timed CPU baseline and the checker."""
import ctypes as C
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import deepreadmapper_b200 as drm
    from deepreadmapper_b200 import synth
    from oracle import bindings as ob
    n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    g = synth.random_genome(4_641_652, 42)
    reads, _ = synth.simulate_reads(g, n_reads, 150, seed=7)
    qual = b"I" * 150
    fastq = b"".join(b"@read%d/1\n" % i + bytes(r) + b"\n+\n" + qual + b"\n" for i, r in enumerate(reads))
    fasta = b">chr synthetic\n" + b"\n".join(g[i:i + 80] for i in range(0, len(g), 80)) + b"\n"
    out = {"reads": n_reads, "fastq_bytes": len(fastq), "fasta_bytes": len(fasta)}
    os.environ["DRM_INPUT_VERBOSE"] = "1"
    for rep in range(3):
        t0 = time.perf_counter()
        rd = drm.Reads(fastq)
        out["fastq_device_wall_ms"] = (time.perf_counter() - t0) * 1e3
        assert rd.n == n_reads
        if rep < 2:
            rd.close()
    for rep in range(3):
        t0 = time.perf_counter()
        clean = drm.fasta_clean(fasta)
        out["fasta_device_wall_ms"] = (time.perf_counter() - t0) * 1e3
    assert clean == g
    seqs = rd.sequences()
    assert seqs[0] == b"<" + bytes(reads[0]) + b">" and seqs[-1] == b"<" + bytes(reads[-1]) + b">"
    if ob.have_ref():
        R = ob.ref()
        with tempfile.TemporaryDirectory() as td:
            pq, pa = os.path.join(td, "r.fastq"), os.path.join(td, "g.fa")
            open(pq, "wb").write(fastq)
            open(pa, "wb").write(fasta)
            n, sb, ib = C.c_int64(0), C.c_int64(0), C.c_int64(0)
            t0 = time.perf_counter()
            assert R.ref_preprocess_fastq(pq.encode(), None, None, None, None, C.byref(n), C.byref(sb), C.byref(ib)) == 0
            out["fastq_reference_host_ms"] = (time.perf_counter() - t0) * 1e3
            assert n.value == n_reads and sb.value == rd.nbytes
            buf = C.create_string_buffer(len(fasta))
            t0 = time.perf_counter()
            ln = R.ref_extract_fasta_sequence(pa.encode(), buf, len(fasta))
            out["fasta_reference_host_ms"] = (time.perf_counter() - t0) * 1e3
            assert buf.raw[:ln] == g
    print(json.dumps(out))


if __name__ == "__main__":
    main()
