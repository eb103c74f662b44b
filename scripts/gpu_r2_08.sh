# round 2, GPU call 8: tcgen05 rerank with per-K-block accumulators: tests, timings, ncu capture (tensor pipe)
set -x
mkdir -p gpurun_out
timeout -s KILL 300 python -m pytest tests/test_gpu_round2.py -q -m gpu --timeout 120 -k "tensor_rerank" > gpurun_out/r2_08_pytest_tc.log 2>&1; echo "rc=$?"; tail -n 12 gpurun_out/r2_08_pytest_tc.log | cut -c1-300
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout -s KILL 400 $SB --dims 768 --metric COSINE --route ternary --L 80 256 --flags 32 --tag rerank_tc > gpurun_out/r2_08_rerank_tc.json 2> gpurun_out/r2_08_rerank_tc.err
timeout -s KILL 400 $SB --dims 128 --metric L2 --route int8 --L 56 128 --flags 32 --tag rerank_tc_d128 > gpurun_out/r2_08_rerank_tc_d128.json 2> gpurun_out/r2_08_rerank_tc_d128.err
tail -n 2 gpurun_out/r2_08_rerank_*.err | cut -c1-600
NCU="ncu --set full --clock-control none --import-source on"
timeout -s KILL 600 $NCU -k regex:rerank_tc_kernel -s 1 -c 1 -o gpurun_out/r2_08_prof_rerank_tc python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 2 --warmup 1 --dims 768 --metric COSINE --route ternary --L 80 --flags 32 > gpurun_out/r2_08_ncu_rerank_tc.log 2>&1
ls -la gpurun_out/r2_08*.ncu-rep
timeout 900 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/r2_08_pytest.log 2>&1; tail -n 6 gpurun_out/r2_08_pytest.log
