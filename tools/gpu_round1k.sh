#!/bin/bash
# 2-GPU session: the driver's N=2 launch of the dense bench, and the position-sharded workload (all-gather + merge).
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.log; echo "n2 rc=$?"
cat gpurun_out/bench_n2.json; tail -4 gpurun_out/bench_n2.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload sharded --steps 3 --warmup 2 > gpurun_out/bench_sharded_n2.json 2> gpurun_out/bench_sharded_n2.log; echo "sharded rc=$?"
cat gpurun_out/bench_sharded_n2.json; tail -4 gpurun_out/bench_sharded_n2.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.log; cat gpurun_out/bench_ref.json
