# exact search without the visited cache at 6 CTAs per SM (80 registers, 9.2 KB of shared memory per warp) against the default (cache, 5 CTAs)
for v in 512 0 512 0; do
  DRM_SEARCH_VIS=$v timeout 300 python bench.py --steps 3 --warmup 3 --skip-cpu --no-realistic --reads 262144 > $O/bench_vis_$v.json 2> $O/bench_vis_$v.log
  show vis_$v $O/bench_vis_$v.json
done
