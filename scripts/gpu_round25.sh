set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > gpurun_out/pytest_gpu_r25.log
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --route ternary --nq 10000 --L 64 80 96 128 192 --steps 10"
timeout 200 $SB --tag default 2>&1 | grep "'L'" | sed "s/^/default /" >> gpurun_out/tune_ternary_r25.log
timeout 200 $SB --tune filter_log2=12 --tag lg12 2>&1 | grep "'L'" | sed "s/^/lg12 /" >> gpurun_out/tune_ternary_r25.log
timeout 200 $SB --tune filter_log2=11 --tag lg11 2>&1 | grep "'L'" | sed "s/^/lg11 /" >> gpurun_out/tune_ternary_r25.log
python bench.py --steps 50 --warmup 5 --no-cpu > gpurun_out/bench_r25.json 2> gpurun_out/bench_r25.err
