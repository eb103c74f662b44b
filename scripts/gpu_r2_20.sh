# round 2, GPU call 20: GPU suite and the N = 1 bench line with the direct-mapped filter on the sidecar routes
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/r2_20_pytest.log 2>&1; tail -n 6 gpurun_out/r2_20_pytest.log | cut -c1-300
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_20_bench.json 2> gpurun_out/r2_20_bench.err; grep "single query" gpurun_out/r2_20_bench.err; cut -c1-400 gpurun_out/r2_20_bench.json
