# L2 rerank: parity (batched, thread-per-query selection) + size-realistic bench; embedder tests with prefetch default; stream-overlap probe
timeout 900 python -m pytest tests/test_gpu_zz_l2.py tests/test_gpu_zz_embed.py -x -q > $O/pytest_l2_embed.log 2>&1; echo "pytest rc=$?" >> $O/pytest_l2_embed.log; tail -5 $O/pytest_l2_embed.log
DRM_L2_BATCH_CANDS=100 timeout 600 python tests/l2_gpu_check.py 23 > $O/l2_check_smallbatch.log 2>&1; echo "small-batch rc=$?" >> $O/l2_check_smallbatch.log; tail -3 $O/l2_check_smallbatch.log
timeout 900 python tools/l2_bench.py > $O/l2_bench.log 2>&1; echo "l2_bench rc=$?"; tail -4 $O/l2_bench.log | cut -c1-600
timeout 600 python tools/overlap_probe.py > $O/overlap_probe.log 2>&1; echo "overlap rc=$?"; tail -12 $O/overlap_probe.log
