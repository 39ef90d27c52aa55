#!/bin/bash
# SW kernel A/B: the add as an fp32 FADD (either fma pipe half) instead of HADD2 (fmaheavy only), with the clamp mixes.
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -k "f32 and (rerank or pipeline)" > gpurun_out/pytest_f32.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_f32.log
tail -5 gpurun_out/pytest_f32.log
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], d["stages_ms"], "sw_gcups", round(d["sw_gcups"], 1), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]))
except Exception as e:
    print(sys.argv[1], "failed", e)
PY
}
for v in blend2 f32p16 f32b2 f32b3 f32b4; do
  DRM_SW_KERNEL=$v python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_sw_$v.json 2> gpurun_out/bench_sw_$v.log
  show $v gpurun_out/bench_sw_$v.json
done
