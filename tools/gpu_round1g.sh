#!/bin/bash
set -x
mkdir -p gpurun_out
./build/ubench_pipes > gpurun_out/ubench_pipes.txt 2>&1
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
cat gpurun_out/ubench_pipes.txt
