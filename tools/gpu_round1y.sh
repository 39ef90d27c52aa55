#!/bin/bash
# Pair-expansion search kernel: parity and A/B.
set -x
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -k "pair" > gpurun_out/pytest_pair.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_pair.log
tail -12 gpurun_out/pytest_pair.log
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], d["stages_ms"], "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["kernel_counters_per_read"])
except Exception as e:
    print(sys.argv[1], "failed", e)
PY
}
for v in pair reg; do
  DRM_SEARCH_KERNEL=$v timeout 200 python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_search_$v.json 2> gpurun_out/bench_search_$v.log
  show search_$v gpurun_out/bench_search_$v.json
done
