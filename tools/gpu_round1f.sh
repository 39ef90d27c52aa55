#!/bin/bash
mkdir -p gpurun_out
for i in 1 2 3 4; do python tools/debug_sw_flaky.py > gpurun_out/flaky_$i.txt 2>&1; grep -c "rows: \[[0-9]" gpurun_out/flaky_$i.txt; done
for i in 1 2 3; do python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "rerank_random" 2>&1 | tail -2; done
