# parity of every SW kernel variant, then A/B of the row-loop forms of the profile kernel (offset staging, rows per trip)
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "rerank or sw" > $O/pytest_sw.log 2>&1; echo "pytest rc=$?" >> $O/pytest_sw.log
tail -4 $O/pytest_sw.log
for v in blend4 blend4u8 blend3o8 blend3o8; do
  DRM_SW_KERNEL=$v timeout 300 python bench.py --steps 3 --warmup 3 --skip-cpu --no-realistic --reads 262144 > $O/bench_sw_$v.json 2> $O/bench_sw_$v.log
  show sw_$v $O/bench_sw_$v.json
done
