# tensor-core embedder, layer 2 with the products issued by a third warpgroup (setmaxnreg; DRM_EMBED_TC_WS=1) against the default:
# parity first (the tensor-core legs of the embedder tests under the switch), then throughput of each form
DRM_EMBED_TC_WS=1 timeout 600 python -m pytest tests/test_gpu_zz_embed.py -x -q > $O/pytest_embed_ws.log 2>&1; echo "pytest rc=$?" >> $O/pytest_embed_ws.log; tail -3 $O/pytest_embed_ws.log
for ws in 0 1 0 1; do
  DRM_EMBED_TC_WS=$ws timeout 300 python tools/embed_bench.py --out $O/embed_bench_ws$ws.json > $O/embed_bench_ws$ws.out 2> $O/embed_bench_ws$ws.log
  python -c "import json;d=json.load(open('$O/embed_bench_ws$ws.json'));print('WS=$ws tc reads/s %.3e windows/s %.3e maxdiff %.2e'%(d['tc_reads_per_s'],d['tc_windows_per_s'],d['tc_max_abs_diff_vs_fp32']))"
done
