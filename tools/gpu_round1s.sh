#!/bin/bash
# Round-end check: full GPU test suite, smoke(), default bench, reference arm.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.log; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench.json'))
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), d['stages_ms'], 'roofline', d['roofline']['bound'], round(d['roofline']['frac'],3), d['roofline']['hbm_view'], 'search frac', round(d['roofline_search']['frac'],4), 'cpu', round(d['cpu_baseline']['value']))"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.log; echo "ref rc=$?"; head -c 400 gpurun_out/bench_ref.json
