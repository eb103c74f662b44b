# round 2, GPU call 42: verification of the committed state: GPU suite, smoke, c4 with the code sidecar, bench.py (no flags)
set -x
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r2_42_pytest.log 2>&1; tail -n 3 gpurun_out/r2_42_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_42_smoke.log 2>&1; tail -n 1 gpurun_out/r2_42_smoke.log
timeout 1200 python bench.py > gpurun_out/r2_42_bench.json 2> gpurun_out/r2_42_bench.err; tail -c 300 gpurun_out/r2_42_bench.json
timeout 600 python bench.py --impl reference > gpurun_out/r2_42_bench_ref.json 2> gpurun_out/r2_42_bench_ref.err; tail -c 400 gpurun_out/r2_42_bench_ref.json
