# round 49: 2 x B200: the NCCL paths of tests/test_sharded.py on the final code, bench.py --gpus 2
set -x
timeout 300 python -m pytest tests/test_sharded.py -x -q -m gpu > gpurun_out/pytest_sharded_n2_r49.log 2>&1; tail -n 3 gpurun_out/pytest_sharded_n2_r49.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/bench_n2_r49.json 2> gpurun_out/bench_n2_r49.err; tail -n 1 gpurun_out/bench_n2_r49.json | cut -c1-400
