# L2 rerank stage with the window-embedding cache: parity (cache forced / default / off), then the stage bench with and without it
timeout 900 python -m pytest tests/test_gpu_zz_l2.py tests/test_gpu_zz_embed.py -x -q > $O/pytest_l2.log 2>&1; echo "pytest rc=$?" >> $O/pytest_l2.log; tail -4 $O/pytest_l2.log
timeout 900 python tools/l2_bench.py --out $O/l2_bench_cache.json > $O/l2_bench_cache.out 2> $O/l2_bench_cache.log; tail -2 $O/l2_bench_cache.log; head -c 1500 $O/l2_bench_cache.json
DRM_L2_CACHE_GB=0 timeout 900 python tools/l2_bench.py --out $O/l2_bench_nocache.json > $O/l2_bench_nocache.out 2> $O/l2_bench_nocache.log; head -c 700 $O/l2_bench_nocache.json
