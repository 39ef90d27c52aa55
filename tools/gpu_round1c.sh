#!/bin/bash
set -x
mkdir -p gpurun_out
python tools/debug_sw_variants.py > gpurun_out/debug_sw.txt 2>&1
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for v in smem reg; do
  DRM_SEARCH_KERNEL=$v python bench.py --steps 3 --warmup 2 --skip-cpu --reads 262144 > gpurun_out/bench_search_$v.json 2> gpurun_out/bench_search_$v.log
  python - <<PY
import json
d=json.load(open("gpurun_out/bench_search_$v.json"))
print("$v", d["stages_ms"], "value", round(d["value"]), d["kernel_counters_per_read"])
PY
done
ncu --set full --clock-control none --import-source on -k regex:hnswpq_search_reg_kernel -c 1 -o gpurun_out/search_reg_full -f \
    python bench.py --steps 1 --warmup 1 --skip-cpu --reads 131072 > gpurun_out/ncu_search.log 2>&1
