# round 2, GPU call 6 (2 GPUs): the GPU suite (device groups on two real devices, NCCL paths), bench.py under torchrun at N = 2
set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/r2_06_pytest.log 2>&1; tail -n 12 gpurun_out/r2_06_pytest.log
cp gpurun_out/concurrent_point_queries.json gpurun_out/r2_06_concurrent_point_queries.json
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_06_bench_n2.json 2> gpurun_out/r2_06_bench_n2.err; tail -n 4 gpurun_out/r2_06_bench_n2.err | cut -c1-400; cut -c1-600 gpurun_out/r2_06_bench_n2.json
