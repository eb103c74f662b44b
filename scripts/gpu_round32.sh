set -x
N=$(nvidia-smi -L | wc -l)
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 5 --cpu-seconds 5 > gpurun_out/bench_n${N}_r32.json 2> gpurun_out/bench_n${N}_r32.err
