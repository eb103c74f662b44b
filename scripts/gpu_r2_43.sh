# round 2, GPU call 43: one-warp CTAs only for batches of a few rounds: c4 (100 k queries) and the headline again
set -x
mkdir -p gpurun_out
timeout 1200 python bench.py > gpurun_out/r2_43_bench.json 2> gpurun_out/r2_43_bench.err; tail -c 300 gpurun_out/r2_43_bench.json
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py tests/test_zz_no_alloc.py -x -q -m gpu --timeout 900 > gpurun_out/r2_43_pytest.log 2>&1; tail -n 2 gpurun_out/r2_43_pytest.log
