#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
for v in blend2 r2a0 r2a1 r2a2 r2a4; do
  DRM_SW_KERNEL=$v python bench.py --steps 3 --warmup 2 --skip-cpu --reads 262144 > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.log
  python - <<PY
import json
d=json.load(open("gpurun_out/bench_$v.json"))
print("$v", d["stages_ms"], "sw_gcups", round(d["sw_gcups"],1), "value", round(d["value"]))
PY
done
DRM_SW_KERNEL=r2a1 ncu --set full --clock-control none --import-source on -k regex:sw_rerank_profile_kernel -c 1 -o gpurun_out/sw_r2a1_full -f \
    python bench.py --steps 1 --warmup 1 --skip-cpu --reads 131072 > gpurun_out/ncu_sw.log 2>&1
