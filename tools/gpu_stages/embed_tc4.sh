# tensor-core embedder (final form) + the L2 stage on top of it
timeout 600 python -m pytest tests/test_gpu_zz_embed.py tests/test_gpu_zz_l2.py -x -q -s > $O/pytest_embed.log 2>&1; echo "pytest rc=$?" >> $O/pytest_embed.log
tail -4 $O/pytest_embed.log
timeout 300 python tools/embed_bench.py --out $O/embed_bench_tc.json > $O/embed_bench_tc.out 2> $O/embed_bench_tc.log; cat $O/embed_bench_tc.out
timeout 900 python tools/l2_bench.py > $O/l2_bench.json 2> $O/l2_bench.log; tail -3 $O/l2_bench.log; cat $O/l2_bench.json
