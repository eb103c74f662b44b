# round 2, GPU call 47 (N GPUs from $NG): bench.py under torchrun with the final kernels
set -x
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $NG --steps 20 --warmup 5 > gpurun_out/r2_47_bench_n$NG.json 2> gpurun_out/r2_47_bench_n$NG.err; tail -n 2 gpurun_out/r2_47_bench_n$NG.err | cut -c1-300; cut -c1-400 gpurun_out/r2_47_bench_n$NG.json
