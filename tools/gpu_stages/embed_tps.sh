# threads per sequence in layer 2's gate phase (2 against 4) on the current gate code
for tps in 22 24 22 24; do
  DRM_EMBED_TC_TPS=$tps timeout 300 python tools/embed_bench.py --out $O/embed_bench_tps$tps.json > $O/embed_bench_tps$tps.out 2> $O/embed_bench_tps$tps.log
  python -c "import json;d=json.load(open('$O/embed_bench_tps$tps.json'));print('TPS=$tps tc reads/s %.3e windows/s %.3e maxdiff %.2e'%(d['tc_reads_per_s'],d['tc_windows_per_s'],d['tc_max_abs_diff_vs_fp32']))"
done
