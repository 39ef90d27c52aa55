#!/bin/bash
# 2-GPU session: the driver's N=2 launch of the dense bench; the position-sharded workload (all-gather + merge) at the E. coli
# size and with two config-5-sized shards (2 x 62.5 M vectors).
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.log; echo "n2 rc=$?"
cat gpurun_out/bench_n2.json; grep "^\[bench\]" gpurun_out/bench_n2.log | tail -6
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload sharded --steps 3 --warmup 2 > gpurun_out/bench_sharded_n2.json 2> gpurun_out/bench_sharded_n2.log; echo "sharded rc=$?"
cat gpurun_out/bench_sharded_n2.json; grep "^\[bench\]" gpurun_out/bench_sharded_n2.log | tail -4
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --workload sharded --genome-bases 62500000 --reads 1000000 --steps 2 --warmup 1 > gpurun_out/bench_sharded125m_n2.json 2> gpurun_out/bench_sharded125m_n2.log; echo "sharded125m rc=$?"
cat gpurun_out/bench_sharded125m_n2.json; grep "^\[bench\]" gpurun_out/bench_sharded125m_n2.log | tail -4
