#!/bin/bash
# GPU session: parity tests, full bench (both arms), launch list, ncu full captures of the two hot kernels, SW variant A/B.
set -x
mkdir -p gpurun_out
nproc > gpurun_out/nproc.txt; nvidia-smi > gpurun_out/nvidia_smi.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.log; echo "bench rc=$?" >> gpurun_out/bench.log
cat gpurun_out/bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.log
for v in profile blend2 blend4 r2a2 r2a4; do
  DRM_SW_KERNEL=$v python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.log
  python - <<PY
import json
d=json.load(open("gpurun_out/bench_$v.json"))
print("$v", d["stages_ms"], "sw_gcups", round(d["sw_gcups"],1), "value", round(d["value"]))
PY
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --skip-cpu > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:hnswpq_search_reg_kernel -c 1 -o gpurun_out/search_reg_full -f \
    python bench.py --steps 1 --warmup 1 --skip-cpu --reads 131072 > gpurun_out/ncu_search.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sw_rerank_profile_kernel -c 1 -o gpurun_out/sw_profile_full -f \
    python bench.py --steps 1 --warmup 1 --skip-cpu --reads 131072 > gpurun_out/ncu_sw.log 2>&1
ls -la gpurun_out
