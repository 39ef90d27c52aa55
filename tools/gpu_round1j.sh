#!/bin/bash
# GPU session: parity with the visited cache, A/B, EF/K sweep (configs[3]), sparse workload (configs[2]), ncu v3 captures.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.log; echo "bench rc=$?" >> gpurun_out/bench.log
tail -3 gpurun_out/bench.log
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], d["stages_ms"], "sw_gcups", round(d["sw_gcups"], 1), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["kernel_counters_per_read"])
except Exception as e:
    print(sys.argv[1], "failed", e)
PY
}
show default gpurun_out/bench.json
DRM_SEARCH_VIS=0 python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_novis.json 2> gpurun_out/bench_novis.log
show novis gpurun_out/bench_novis.json
DRM_SEARCH_VIS=1 python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_vis.json 2> gpurun_out/bench_vis.log
show vis gpurun_out/bench_vis.json
python bench.py --workload sparse --steps 3 --warmup 2 > gpurun_out/bench_sparse.json 2> gpurun_out/bench_sparse.log; echo "sparse rc=$?"
cat gpurun_out/bench_sparse.json; tail -3 gpurun_out/bench_sparse.log
python tools/sweep_ef_k.py --out gpurun_out/sweep_ef_k.json > gpurun_out/sweep_ef_k.jsonl 2> gpurun_out/sweep.log; echo "sweep rc=$?"
tail -3 gpurun_out/sweep.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --skip-cpu > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:hnswpq_search_reg_kernel -c 1 -o gpurun_out/search_reg_full -f \
    python bench.py --steps 1 --warmup 1 --skip-cpu --reads 131072 > gpurun_out/ncu_search.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sw_rerank_profile_kernel -c 1 -o gpurun_out/sw_profile_full -f \
    python bench.py --steps 1 --warmup 1 --skip-cpu --reads 131072 > gpurun_out/ncu_sw.log 2>&1
ls gpurun_out
