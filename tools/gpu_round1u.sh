#!/bin/bash
# SW kernel A/B: 80-register build (3 CTAs of 256 threads per SM instead of 2).
set -x
mkdir -p gpurun_out
DRM_SW_DENSE=1 python -m pytest tests/test_gpu_parity.py -x -q -k "rerank and (default or blend2 or r2a2 or profile16)" > gpurun_out/pytest_dense.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dense.log
tail -4 gpurun_out/pytest_dense.log
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], d["stages_ms"], "sw_gcups", round(d["sw_gcups"], 1), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]))
except Exception as e:
    print(sys.argv[1], "failed", e)
PY
}
for v in blend2 profile16 blend4 r2a2 r2a4 f32b3; do
  DRM_SW_DENSE=1 DRM_SW_KERNEL=$v python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_dense_$v.json 2> gpurun_out/bench_dense_$v.log
  show dense_$v gpurun_out/bench_dense_$v.json
done
