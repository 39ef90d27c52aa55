#!/bin/bash
# sw_select in shared memory: parity (all rerank tests) and A/B against the thread-local form.
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -k "rerank or pipeline or one_call or cli" > gpurun_out/pytest_sel.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_sel.log
tail -4 gpurun_out/pytest_sel.log
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], d["stages_ms"], "sw_gcups", round(d["sw_gcups"], 1), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]))
except Exception as e:
    print(sys.argv[1], "failed", e)
PY
}
for v in smem local; do
  DRM_SW_SELECT=$v python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_sel_$v.json 2> gpurun_out/bench_sel_$v.log
  show select_$v gpurun_out/bench_sel_$v.json
done
