#!/bin/bash
# GPU session: parity (incl. the one-call test), default bench, search A/B (visited-cache size, L2 prefetch depth), sparse workload.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], d["stages_ms"], "sw_gcups", round(d["sw_gcups"], 1), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["kernel_counters_per_read"])
except Exception as e:
    print(sys.argv[1], "failed", e)
PY
}
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.log; echo "bench rc=$?"
show default gpurun_out/bench.json
for cfg in "1024 1" "1024 2" "512 0"; do set -- $cfg
  DRM_SEARCH_VIS=$1 DRM_SEARCH_PF=$2 python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_vis$1_pf$2.json 2> gpurun_out/bench_vis$1_pf$2.log
  show vis$1_pf$2 gpurun_out/bench_vis$1_pf$2.json
done
python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_base.json 2> gpurun_out/bench_base.log
show base262k gpurun_out/bench_base.json
python bench.py --workload sparse --steps 3 --warmup 2 > gpurun_out/bench_sparse.json 2> gpurun_out/bench_sparse.log; echo "sparse rc=$?"
cat gpurun_out/bench_sparse.json; tail -3 gpurun_out/bench_sparse.log
