# round 2, GPU call 45: verification: GPU suite, smoke, bench.py (no flags; c3 now with the code sidecar too)
set -x
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r2_45_pytest.log 2>&1; tail -n 3 gpurun_out/r2_45_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_45_smoke.log 2>&1; tail -n 1 gpurun_out/r2_45_smoke.log
timeout 1200 python bench.py > gpurun_out/r2_45_bench.json 2> gpurun_out/r2_45_bench.err; tail -c 300 gpurun_out/r2_45_bench.json
