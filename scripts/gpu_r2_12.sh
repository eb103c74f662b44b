# round 2, GPU call 12 (2 GPUs): GPU suite, bench.py N = 2 under torchrun (c5 with NCCL), both arms
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/r2_12_pytest.log 2>&1; tail -n 6 gpurun_out/r2_12_pytest.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_12_bench_n2.json 2> gpurun_out/r2_12_bench_n2.err; tail -n 3 gpurun_out/r2_12_bench_n2.err | cut -c1-900
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --steps 20 --warmup 5 --impl reference > gpurun_out/r2_12_bench_n2_ref.json 2> gpurun_out/r2_12_bench_n2_ref.err; cut -c1-300 gpurun_out/r2_12_bench_n2_ref.json
