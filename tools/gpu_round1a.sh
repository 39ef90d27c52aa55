#!/bin/bash
# GPU session: parity tests, bench, ncu launch list + full captures of the two hot kernels, micro-benchmarks.
set -x
mkdir -p gpurun_out
nproc > gpurun_out/nproc.txt; nvidia-smi > gpurun_out/nvidia_smi.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
./build/ubench_sw_mix > gpurun_out/ubench_sw_mix.txt 2>&1
./build/ubench_alu > gpurun_out/ubench_alu.txt 2>&1
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.log; echo "bench rc=$?" >> gpurun_out/bench.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.log
# launch list of the same command (single metric pass)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --skip-cpu > gpurun_out/ncu_launch.log 2>&1
# full captures, smaller batch (one launch each)
ncu --set full --clock-control none --import-source on -k regex:hnswpq_search_kernel -c 1 -o gpurun_out/search_full -f \
    python bench.py --steps 1 --warmup 1 --skip-cpu --reads 131072 > gpurun_out/ncu_search.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sw_rerank_kernel -c 1 -o gpurun_out/sw_full -f \
    python bench.py --steps 1 --warmup 1 --skip-cpu --reads 131072 > gpurun_out/ncu_sw.log 2>&1
ls -la gpurun_out
