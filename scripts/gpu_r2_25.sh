# round 2, GPU call 25: verification of the committed state - GPU suite, smoke, bench.py with NO flags, reference arm
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/r2_25_pytest.log 2>&1; tail -n 3 gpurun_out/r2_25_pytest.log | cut -c1-200
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2_25_smoke.log 2>&1; tail -n 1 gpurun_out/r2_25_smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/r2_25_bench_ref.json 2> gpurun_out/r2_25_bench_ref.err; cut -c1-300 gpurun_out/r2_25_bench_ref.json
timeout 900 python bench.py > gpurun_out/r2_25_bench.json 2> gpurun_out/r2_25_bench.err; grep "single query" gpurun_out/r2_25_bench.err; cut -c1-600 gpurun_out/r2_25_bench.json
