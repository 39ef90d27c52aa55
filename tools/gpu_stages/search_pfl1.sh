# search parity sweep (all variants incl. the capped grid), then A/B of the accept-time prefetch target (L2 default / L1)
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "search" > $O/pytest_search.log 2>&1; echo "pytest rc=$?" >> $O/pytest_search.log; tail -3 $O/pytest_search.log
for v in 0 1 0 1; do
  DRM_SEARCH_PFL1=$v timeout 300 python bench.py --steps 3 --warmup 3 --skip-cpu --no-realistic --reads 262144 > $O/bench_pfl1_$v.json 2> $O/bench_pfl1_$v.log
  show pfl1_$v $O/bench_pfl1_$v.json
done
