# the tensor-core form of the embedder: parity (both forms), then throughput A/B
timeout 600 python -m pytest tests/test_gpu_zz_embed.py -x -q -s > $O/pytest_embed.log 2>&1; echo "pytest rc=$?" >> $O/pytest_embed.log
tail -15 $O/pytest_embed.log
timeout 300 python tools/embed_bench.py --out $O/embed_bench_tc.json > $O/embed_bench_tc.out 2> $O/embed_bench_tc.log; tail -3 $O/embed_bench_tc.log; cat $O/embed_bench_tc.out
