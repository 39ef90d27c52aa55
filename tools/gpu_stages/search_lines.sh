# source-line attribution of the exact search kernel (instructions and stall samples per CUDA line)
ncu_kernel hnswpq_search r2_ncu_search_lines python bench.py --steps 1 --warmup 0 --skip-cpu --no-realistic --reads 131072
python tools/ncu_lines.py $O/r2_ncu_search_lines.lines.csv 70 > $O/r2_ncu_search_lines.bylines.txt 2>&1
sort -t'%' -k2 $O/r2_ncu_search_lines.bylines.txt > /dev/null
rm -f $O/r2_ncu_search_lines.lines.csv $O/r2_ncu_search_lines.ncu-rep
