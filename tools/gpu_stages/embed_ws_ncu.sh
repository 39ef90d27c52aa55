# one ncu capture of layer 2 in each form (default / warp-specialised), with the per-source-line stall attribution
for ws in 1 0; do
  rx=gru_tc2_ws; [ $ws = 0 ] && rx='gru_tc_kernel<2'
  DRM_EMBED_TC_WS=$ws ncu --set full --clock-control none --import-source on -k regex:"$rx" -c 1 -o $O/ncu_embed_ws$ws -f python tools/embed_bench.py --reads 37888 --reps 1 --out $O/embed_bench_ncu.json > $O/ncu_embed_ws$ws.log 2>&1; echo "ncu rc=$?"
  ncu -i $O/ncu_embed_ws$ws.ncu-rep --page raw --csv > $O/ncu_embed_ws$ws.raw.csv 2>/dev/null
  ncu -i $O/ncu_embed_ws$ws.ncu-rep --page source --csv > $O/ncu_embed_ws$ws.sass.csv 2>/dev/null
  ncu -i $O/ncu_embed_ws$ws.ncu-rep --page source --print-source cuda,sass --csv > $O/ncu_embed_ws$ws.lines.csv 2>/dev/null
  python tools/ncu_summary.py $O/ncu_embed_ws$ws.raw.csv > $O/ncu_embed_ws$ws.summary.txt 2>&1
  python tools/ncu_sass.py $O/ncu_embed_ws$ws.sass.csv 40 >> $O/ncu_embed_ws$ws.summary.txt 2>&1
  rm -f $O/ncu_embed_ws$ws.ncu-rep
done
