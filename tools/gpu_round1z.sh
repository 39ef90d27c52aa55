#!/bin/bash
# Parity of every search kernel variant after the shared insert helper / low-word duplicate test, and the default bench line.
set -x
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -k "search" > gpurun_out/pytest_search.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_search.log
tail -5 gpurun_out/pytest_search.log
timeout 200 python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_final262k.json 2> gpurun_out/bench_final262k.log
python -c "
import json; d=json.load(open('gpurun_out/bench_final262k.json')); print(d['stages_ms'], round(d['value']), round(d['e2e']['value']), d['kernel_counters_per_read'])"
