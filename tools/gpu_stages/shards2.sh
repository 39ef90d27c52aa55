# multi-GPU: sharded tests (one process / rank per process, all exchange modes), then the driver's launch line with the sharded section
N=$(nvidia-smi -L | wc -l)
nvidia-smi topo -m > $O/topo_n$N.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_shards.py -x -q > $O/pytest_shards_n$N.log 2>&1; echo "pytest shards rc=$?" >> $O/pytest_shards_n$N.log; tail -15 $O/pytest_shards_n$N.log
NCCL_DEBUG=WARN timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.log
echo "bench n$N rc=$?"; tail -12 $O/bench_n$N.log; cat $O/bench_n$N.json
