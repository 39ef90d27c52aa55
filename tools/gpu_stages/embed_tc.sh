# tensor-core embedder: parity on the default form (+ the L2 stage on top of it), threads-per-sequence A/B, one ncu capture per kernel
timeout 600 python -m pytest tests/test_gpu_zz_embed.py tests/test_gpu_zz_l2.py -x -q -s > $O/pytest_embed.log 2>&1; echo "pytest rc=$?" >> $O/pytest_embed.log
tail -4 $O/pytest_embed.log
for tps in 22 24; do
  DRM_EMBED_TC_TPS=$tps timeout 300 python tools/embed_bench.py --out $O/embed_bench_tps$tps.json > $O/embed_bench_tps$tps.out 2> $O/embed_bench_tps$tps.log
  python - $tps $O/embed_bench_tps$tps.json <<'PY'
import json, sys
d = json.load(open(sys.argv[2])); print("tps", sys.argv[1], "tc reads/s", round(d["tc_reads_per_s"]), "windows/s", round(d["tc_windows_per_s"]), "maxdiff", d["tc_max_abs_diff_vs_fp32"])
PY
done
ncu --set full --clock-control none --import-source on -k regex:gru_tc -c 2 -o $O/r2_ncu_embed_tc -f python tools/embed_bench.py --reads 37888 --reps 1 --out $O/embed_bench_ncu.json > $O/ncu_embed_tc.log 2>&1; echo "ncu rc=$?"
ncu -i $O/r2_ncu_embed_tc.ncu-rep --page raw --csv > $O/r2_ncu_embed_tc.raw.csv 2>/dev/null
ncu -i $O/r2_ncu_embed_tc.ncu-rep --page source --csv > $O/r2_ncu_embed_tc.sass.csv 2>/dev/null
python tools/ncu_summary.py $O/r2_ncu_embed_tc.raw.csv > $O/r2_ncu_embed_tc.summary.txt 2>&1
rm -f $O/r2_ncu_embed_tc.ncu-rep
