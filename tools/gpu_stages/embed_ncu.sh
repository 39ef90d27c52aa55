# one ncu capture per layer of the tensor-core embedder (default form), summary only
ncu --set full --clock-control none --import-source on -k regex:gru_tc_kernel -c 2 -o $O/ncu_embed_tc -f python tools/embed_bench.py --reads 37888 --reps 1 --out $O/embed_bench_ncu.json > $O/ncu_embed_tc.log 2>&1; echo "ncu rc=$?"
ncu -i $O/ncu_embed_tc.ncu-rep --page raw --csv > $O/ncu_embed_tc.raw.csv 2>/dev/null
python tools/ncu_summary.py $O/ncu_embed_tc.raw.csv > $O/ncu_embed_tc.summary.txt 2>&1
rm -f $O/ncu_embed_tc.ncu-rep
