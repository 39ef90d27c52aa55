#!/bin/bash
# Round-end: full GPU suite, smoke, default bench, reference arm, launch list and ncu captures of the two hot kernels.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.log; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench.json'))
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), d['stages_ms'], 'roofline', round(d['roofline']['frac'],3), 'search frac', round(d['roofline_search']['frac'],4), 'cpu', round(d['cpu_baseline']['value']), d['clocks'])"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.log; echo "ref rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --skip-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:sw_rerank_profile_kernel -c 1 -o gpurun_out/sw_profile_full -f python bench.py --steps 1 --warmup 1 --skip-cpu --reads 131072 > gpurun_out/ncu_sw.log 2>&1; echo "ncu sw rc=$?"
ncu --set full --clock-control none --import-source on -k regex:hnswpq_search_reg_kernel -c 1 -o gpurun_out/search_reg_full -f python bench.py --steps 1 --warmup 1 --skip-cpu --reads 131072 > gpurun_out/ncu_search.log 2>&1; echo "ncu search rc=$?"
