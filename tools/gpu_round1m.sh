#!/bin/bash
# GPU session: GPU index builder - tests, then a full-size (9.28 M vectors, EFC 200) build with graph statistics, search time and recall.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_zbuilder.py -x -q > gpurun_out/pytest_builder.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_builder.log
tail -40 gpurun_out/pytest_builder.log
DRM_BUILD_VERBOSE=1 timeout 900 python tools/build_compare.py --out gpurun_out/build_compare.json > gpurun_out/build_compare.out 2> gpurun_out/build_compare.log; echo "compare rc=$?"
tail -12 gpurun_out/build_compare.log
