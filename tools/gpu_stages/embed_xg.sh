# tensor-core embedder, layer 2: the next step's input products issued in groups between the gates (XG per group) against all
# sixteen before the gates (XG = 16): parity first, then throughput of each form and the phase clocks of the default
timeout 600 python -m pytest tests/test_gpu_zz_embed.py -x -q > $O/pytest_embed_xg.log 2>&1; echo "pytest rc=$?" >> $O/pytest_embed_xg.log; tail -3 $O/pytest_embed_xg.log
for xg in 16 4 2 16 2; do
  DRM_EMBED_TC_XG=$xg timeout 300 python tools/embed_bench.py --out $O/embed_bench_xg$xg.json > $O/embed_bench_xg$xg.out 2> $O/embed_bench_xg$xg.log
  python -c "import json;d=json.load(open('$O/embed_bench_xg$xg.json'));print('XG=$xg tc reads/s %.3e windows/s %.3e maxdiff %.2e'%(d['tc_reads_per_s'],d['tc_windows_per_s'],d['tc_max_abs_diff_vs_fp32']))"
done
DRM_EMBED_TC_PHASES=1 timeout 300 python tools/embed_bench.py --reads 37888 --reps 1 --out $O/embed_phases.json > $O/embed_phases.out 2> $O/embed_phases.log; grep "embed tc" $O/embed_phases.log | tail -2
