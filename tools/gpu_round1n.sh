#!/bin/bash
# GPU session: default bench on the GPU-built index, launch list, ncu captures of the two builder kernels, one config-5-sized shard.
set -x
mkdir -p gpurun_out
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.log; echo "bench rc=$?"
tail -4 gpurun_out/bench.log; cat gpurun_out/bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --skip-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:build_insert_kernel --launch-skip 300 -c 1 -o gpurun_out/build_insert_full -f python tools/build_compare.py --reads 4096 --sample 64 > gpurun_out/ncu_build_insert.log 2>&1; echo "ncu insert rc=$?"
ncu --set full --clock-control none --import-source on -k regex:build_backlink_kernel --launch-skip 300 -c 1 -o gpurun_out/build_backlink_full -f python tools/build_compare.py --reads 4096 --sample 64 > gpurun_out/ncu_build_backlink.log 2>&1; echo "ncu backlink rc=$?"
timeout 900 python bench.py --workload sharded --genome-bases 31250000 --reads 1000000 --steps 2 --warmup 1 > gpurun_out/bench_shard62m.json 2> gpurun_out/bench_shard62m.log; echo "shard rc=$?"
cat gpurun_out/bench_shard62m.json; tail -5 gpurun_out/bench_shard62m.log
