#!/bin/bash
# GPU session: parity after the SW staging / PCPL=38 / search insert / e2e workspace changes, then A/B runs.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
./build/ubench_pipes2 > gpurun_out/ubench_pipes2.txt 2>&1
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.log; echo "bench rc=$?" >> gpurun_out/bench.log
tail -3 gpurun_out/bench.log
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], d["stages_ms"], "sw_gcups", round(d["sw_gcups"], 1), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]))
except Exception as e:
    print(sys.argv[1], "failed", e)
PY
}
show default gpurun_out/bench.json
for v in blend2 r2a2 r2a4; do for t in 256 128; do
  DRM_SW_KERNEL=$v DRM_SW_THREADS=$t python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > gpurun_out/bench_${v}_$t.json 2> gpurun_out/bench_${v}_$t.log
  show ${v}_$t gpurun_out/bench_${v}_$t.json
done; done
for c in 16384 65536 131072; do
  DRM_CHUNK=$c python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/bench_chunk$c.json 2> gpurun_out/bench_chunk$c.log
  show chunk$c gpurun_out/bench_chunk$c.json
done
ls gpurun_out
