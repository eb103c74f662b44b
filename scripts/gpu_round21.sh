set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_gpu_r21.log
timeout 300 python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --route ternary --nq 10000 --L 64 80 96 --steps 20 --tag C3_ternary_batchmerge_1gpu > gpurun_out/shard_c3_ternary_n1_r21.json 2> gpurun_out/shard_c3_ternary_n1_r21.err
timeout 300 python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 128 --metric COSINE --route ternary --nq 10000 --L 64 128 --steps 20 --tag d128_ternary_1gpu > gpurun_out/shard_d128_ternary_n1_r21.json 2> gpurun_out/shard_d128_ternary_n1_r21.err
python bench.py --steps 50 --warmup 5 --no-cpu > gpurun_out/bench_r21.json 2> gpurun_out/bench_r21.err
