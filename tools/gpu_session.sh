#!/bin/bash
# One parameterised GPU session script (replaces the per-lease scratch scripts of round 1).
#   usage (on the box, through gpurun):  bash tools/gpu_session.sh <stage> [<stage> ...]
# Every stage writes under gpurun_out/ ; numbers that are kept get copied into profiles/ by hand afterwards.
set -x
mkdir -p gpurun_out
O=gpurun_out

show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], d.get("stages_ms"), "value", round(d["value"]), "e2e", round(d.get("e2e", {}).get("value", 0)), d.get("kernel_counters_per_read"),
          "sw_frac", round(d.get("roofline", {}).get("frac", 0), 4), "search_frac", round(d.get("roofline_search", {}).get("frac", 0), 4))
except Exception as e:
    print(sys.argv[1], "failed", e)
PY
}

ncu_kernel() {  # ncu_kernel <regex> <outname> <command...>: one --set full capture with source, exported as raw + sass csv
  local rx=$1 name=$2; shift 2
  ncu --set full --clock-control none --import-source on -k regex:$rx -c 1 -o $O/$name -f "$@" > $O/$name.log 2>&1; echo "ncu $name rc=$?"
  ncu -i $O/$name.ncu-rep --page raw --csv > $O/$name.raw.csv 2>/dev/null
  ncu -i $O/$name.ncu-rep --page source --csv > $O/$name.sass.csv 2>/dev/null
  ncu -i $O/$name.ncu-rep --page source --print-source cuda,sass --csv > $O/$name.lines.csv 2>/dev/null
  python tools/ncu_summary.py $O/$name.raw.csv > $O/$name.summary.txt 2>&1
  python tools/ncu_sass.py $O/$name.sass.csv 40 >> $O/$name.summary.txt 2>&1
  ls -la $O/$name.ncu-rep
  [ $(stat -c %s $O/$name.ncu-rep) -gt 20000000 ] && rm -f $O/$name.ncu-rep
}

for stage in "$@"; do
case $stage in
pytest)
  timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -5 $O/pytest_gpu.log ;;
noalloc)
  # alone in a fresh process, and again after another GPU test file in the same process
  timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > $O/pytest_noalloc.log 2>&1; echo "pytest rc=$?" >> $O/pytest_noalloc.log; tail -4 $O/pytest_noalloc.log
  timeout 600 python -m pytest tests/test_gpu_align.py tests/test_zz_no_alloc.py -x -q > $O/pytest_noalloc_after.log 2>&1; echo "pytest rc=$?" >> $O/pytest_noalloc_after.log; tail -4 $O/pytest_noalloc_after.log ;;
pfa_ab)
  # accept-time L2 prefetch: off / every accepted vertex / gated to the lowest 1/2, 1/4, 1/8 of the [popped, root] interval
  timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "search" > $O/pytest_search.log 2>&1; echo "pytest rc=$?" >> $O/pytest_search.log; tail -3 $O/pytest_search.log
  for pfa in 1 2 3 4 0; do
    DRM_SEARCH_PFA=$pfa timeout 300 python bench.py --steps 3 --warmup 3 --skip-cpu --no-realistic --reads 262144 > $O/bench_pfa$pfa.json 2> $O/bench_pfa$pfa.log
    show pfa$pfa $O/bench_pfa$pfa.json
  done
  for pfa in 2 3; do
    DRM_SEARCH_PFA=$pfa ncu_kernel hnswpq_search r2_ncu_search_pfa$pfa python bench.py --steps 1 --warmup 0 --skip-cpu --no-realistic --reads 131072
    grep -E "dram__bytes_read.sum \[|gpu__time_duration" $O/r2_ncu_search_pfa$pfa.summary.txt
  done ;;
bench)
  timeout 900 python bench.py > $O/bench.json 2> $O/bench.log; echo "bench rc=$?"; show bench $O/bench.json ;;
bench_ref)
  timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.log; echo "ref rc=$?"; cat $O/bench_reference.json ;;
bench262)
  timeout 400 python bench.py --steps 3 --warmup 3 --skip-cpu --reads 262144 > $O/bench262.json 2> $O/bench262.log; show bench262 $O/bench262.json ;;
launches)
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 1 --skip-cpu > $O/ncu_launch.log 2>&1; echo "ncu launches rc=$?" ;;
sanitize)
  for tool in memcheck racecheck; do
    for part in search rerank build merge align input; do
      timeout 600 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_case.py --part $part > $O/sanitizer_${tool}_${part}.txt 2>&1
      echo "sanitizer $tool $part rc=$?" | tee -a $O/sanitizer_${tool}_${part}.txt
      grep -E "ERROR SUMMARY|RACECHECK SUMMARY|ok$| ok" $O/sanitizer_${tool}_${part}.txt | tail -4
    done
  done ;;
ncu_search)
  ncu_kernel hnswpq_search r2_ncu_search python bench.py --steps 1 --warmup 0 --skip-cpu --reads 131072 ;;
ncu_sw)
  ncu_kernel sw_rerank_profile r2_ncu_sw python bench.py --steps 1 --warmup 0 --skip-cpu --reads 131072 ;;
embed_ab)
  timeout 300 python tools/embed_bench.py > $O/embed_bench.json 2> $O/embed_bench.log; tail -3 $O/embed_bench.log; cat $O/embed_bench.json ;;
overlap)
  timeout 600 python tools/overlap_probe.py > $O/overlap_probe.log 2>&1; tail -12 $O/overlap_probe.log ;;
l2bench)
  timeout 600 python tools/l2_bench.py > $O/l2_bench.json 2> $O/l2_bench.log; tail -3 $O/l2_bench.log; cat $O/l2_bench.json ;;
sharded)
  # N is taken from the visible GPUs; the driver's launch line
  N=$(nvidia-smi -L | wc -l)
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.log
  echo "bench n$N rc=$?"; tail -5 $O/bench_n$N.log; cat $O/bench_n$N.json ;;
chr8)
  N=$(nvidia-smi -L | wc -l)
  DRM_BENCH_CHR=1 timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 1 --workload sharded > $O/bench_chr$N.json 2> $O/bench_chr$N.log
  echo "chr rc=$?"; tail -8 $O/bench_chr$N.log; cat $O/bench_chr$N.json ;;
*)
  # anything else: a shell snippet file under tools/gpu_stages/<stage>.sh
  if [ -f tools/gpu_stages/$stage.sh ]; then source tools/gpu_stages/$stage.sh; else echo "unknown stage $stage"; fi ;;
esac
done
exit 0
