# tensor-core embedder: parity + throughput + phase clocks (no ncu)
timeout 600 python -m pytest tests/test_gpu_zz_embed.py tests/test_gpu_zz_l2.py -x -q > $O/pytest_embed.log 2>&1; echo "pytest rc=$?" >> $O/pytest_embed.log; tail -3 $O/pytest_embed.log
timeout 300 python tools/embed_bench.py --out $O/embed_bench_tc.json > $O/embed_bench_tc.out 2> $O/embed_bench_tc.log; cat $O/embed_bench_tc.out
DRM_EMBED_TC_PHASES=1 timeout 300 python tools/embed_bench.py --reads 37888 --reps 1 --out $O/embed_phases.json > $O/embed_phases.out 2> $O/embed_phases.log; grep "embed tc" $O/embed_phases.log | tail -2
