set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_gpu_r18.log
timeout 300 python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --nq 10000 --L 32 48 64 --steps 20 --tag C3_exact_G6_1gpu > gpurun_out/shard_c3_exact_n1_r18.json 2> gpurun_out/shard_c3_exact_n1_r18.err
timeout 300 python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --route ternary --nq 10000 --L 64 80 96 --steps 20 --tag C3_ternary_G6_1gpu > gpurun_out/shard_c3_ternary_n1_r18.json 2> gpurun_out/shard_c3_ternary_n1_r18.err
python bench.py --steps 50 --warmup 5 --cpu-seconds 5 > gpurun_out/bench_r18.json 2> gpurun_out/bench_r18.err
