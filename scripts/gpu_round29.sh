# round 29: 8 x B200 again with the current kernels - headline, C3 (exact / ternary), C4, C5 (single-stream sharded search)
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 8 --steps 50 --warmup 5 --cpu-seconds 5 > gpurun_out/bench_n8_r29.json 2> gpurun_out/bench_n8_r29.err
timeout 400 $TR scripts/shard_bench.py --mode sharded --rows 100000000 --dims 96 --nq 10000 --L 48 64 96 --steps 20 --tag C5 > gpurun_out/shard_c5_n8_r29.json 2> gpurun_out/shard_c5_n8_r29.err
timeout 300 $TR scripts/shard_bench.py --mode replicated --rows 10000000 --dims 96 --nq 12500 --L 48 64 96 --steps 20 --tag C4 > gpurun_out/shard_c4_n8_r29.json 2> gpurun_out/shard_c4_n8_r29.err
timeout 300 $TR scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --nq 10000 --L 32 48 --steps 20 --tag C3_exact > gpurun_out/shard_c3_exact_n8_r29.json 2> gpurun_out/shard_c3_exact_n8_r29.err
timeout 300 $TR scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --route ternary --nq 10000 --L 80 96 --steps 20 --tag C3_ternary > gpurun_out/shard_c3_ternary_n8_r29.json 2> gpurun_out/shard_c3_ternary_n8_r29.err
