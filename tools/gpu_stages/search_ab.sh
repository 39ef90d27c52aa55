# parity of every search kernel variant, then A/B of the exact kernel's options against the canonical register-pool kernel
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "search" > $O/pytest_search.log 2>&1; echo "pytest rc=$?" >> $O/pytest_search.log
tail -4 $O/pytest_search.log
for v in "exact 0 1 0" "exact 0 1 1" "exact 0 0 1" "reg 1 0 0"; do
  set -- $v
  DRM_SEARCH_KERNEL=$1 DRM_SEARCH_RA=$2 DRM_SEARCH_PFA=$3 DRM_SEARCH_TMA=$4 timeout 300 python bench.py --steps 3 --warmup 3 --skip-cpu --no-realistic --reads 262144 > $O/bench_search_$1_ra$2_pfa$3_tma$4.json 2> $O/bench_search_$1_ra$2_pfa$3_tma$4.log
  show search_$1_ra$2_pfa$3_tma$4 $O/bench_search_$1_ra$2_pfa$3_tma$4.json
done
