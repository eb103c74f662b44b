# round 2, GPU call 2: NaN debug, the GPU suite after the strict / tie-slot / test fixes + lmd_group, bench.py with other_configs
set -x
mkdir -p gpurun_out
timeout 120 python scripts/debug/special_values.py > gpurun_out/r2_02_special.log 2>&1; tail -n 30 gpurun_out/r2_02_special.log
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 -x --deselect tests/test_gpu_round2.py::test_special_values > gpurun_out/r2_02_pytest.log 2>&1; tail -n 15 gpurun_out/r2_02_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_02_bench.json 2> gpurun_out/r2_02_bench.err; tail -n 5 gpurun_out/r2_02_bench.err; cat gpurun_out/r2_02_bench.json | cut -c1-3000
