# round 2, GPU call 7: the tcgen05 rerank (first run: under a hard timeout), then rerank timings FMA vs tensor
set -x
mkdir -p gpurun_out
timeout -s KILL 300 python -m pytest tests/test_gpu_round2.py -q -m gpu --timeout 120 -k "tensor_rerank" -x > gpurun_out/r2_07_pytest_tc.log 2>&1; echo "rc=$?"; tail -n 30 gpurun_out/r2_07_pytest_tc.log
nvidia-smi --query-gpu=memory.used,utilization.gpu --format=csv
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout -s KILL 400 $SB --dims 768 --metric COSINE --route ternary --L 80 256 --tag rerank_fma > gpurun_out/r2_07_rerank_fma.json 2> gpurun_out/r2_07_rerank_fma.err
timeout -s KILL 400 $SB --dims 768 --metric COSINE --route ternary --L 80 256 --flags 32 --tag rerank_tc > gpurun_out/r2_07_rerank_tc.json 2> gpurun_out/r2_07_rerank_tc.err
tail -n 2 gpurun_out/r2_07_rerank_*.err | cut -c1-700
