# round 2, GPU call 13 (8 GPUs): bench.py under torchrun at N = 8 (headline + other_configs incl. c5 = 100 M rows), single-process group
set -x
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_13_bench_n8.json 2> gpurun_out/r2_13_bench_n8.err; tail -n 3 gpurun_out/r2_13_bench_n8.err | cut -c1-1200
timeout 600 python bench.py --gpus 8 --single-process --steps 20 --warmup 5 > gpurun_out/r2_13_single_process_n8.json 2> gpurun_out/r2_13_single_process_n8.err; cat gpurun_out/r2_13_single_process_n8.json | cut -c1-800; tail -n 3 gpurun_out/r2_13_single_process_n8.err
