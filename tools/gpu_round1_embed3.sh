#!/bin/bash
# One ncu --set full capture of the three embedder kernels (first pass of embed_bench at one full scratch pass).
set -x
mkdir -p gpurun_out
timeout 60 ncu --set full --import-source on --clock-control none -k regex:'gru_kernel|proj2_kernel' -c 3 -f -o gpurun_out/r1_embed_v1 \
    python tools/embed_bench.py --reads 9472 --reps 1 --out gpurun_out/embed_bench_ncu.json > gpurun_out/ncu_embed.log 2>&1
echo "ncu rc=$?"
tail -5 gpurun_out/ncu_embed.log
ls -la gpurun_out/*.ncu-rep
