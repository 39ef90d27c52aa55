# one-GPU sanity of the sharded data plane (W = 1 paths), the whole -m gpu suite, and the sharded bench arm at N = 1
timeout 600 python -m pytest tests/test_gpu_shards.py -x -q > $O/pytest_shards.log 2>&1; echo "pytest shards rc=$?" >> $O/pytest_shards.log; tail -15 $O/pytest_shards.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -8 $O/pytest_gpu.log
timeout 600 python bench.py --workload sharded --reads 262144 --steps 2 --warmup 2 > $O/bench_sharded_n1.json 2> $O/bench_sharded_n1.log; echo "sharded n1 rc=$?"; tail -5 $O/bench_sharded_n1.log; cat $O/bench_sharded_n1.json
