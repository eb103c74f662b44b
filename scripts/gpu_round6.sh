set -x
python -m pytest tests -m gpu -q 2>&1 | tail -12 > gpurun_out/pytest_gpu_r06.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 --cpu-seconds 5 > gpurun_out/bench_n2_r06.json 2> gpurun_out/bench_n2_r06.err
python bench.py --gpus 1 --steps 50 --warmup 5 --cpu-seconds 5 > gpurun_out/bench_n1_r06.json 2> gpurun_out/bench_n1_r06.err
