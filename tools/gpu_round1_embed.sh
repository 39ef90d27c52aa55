#!/bin/bash
# Embedder: GPU parity tests + throughput.
set -x
mkdir -p gpurun_out
timeout 70 python -m pytest tests/test_gpu_zz_embed.py -q > gpurun_out/pytest_embed.log 2>&1
echo "pytest rc=$?"
tail -40 gpurun_out/pytest_embed.log
timeout 40 python tools/embed_bench.py > gpurun_out/embed_bench.log 2>&1
echo "bench rc=$?"
tail -3 gpurun_out/embed_bench.log
