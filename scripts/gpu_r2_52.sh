# round 2, GPU call 52: last sanity of the committed build: GPU suite (-x) and smoke
set -x
mkdir -p gpurun_out
timeout 200 python -m pytest tests -x -q -m gpu --timeout 180 > gpurun_out/r2_52_pytest.log 2>&1; tail -n 2 gpurun_out/r2_52_pytest.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_52_smoke.log 2>&1; tail -n 1 gpurun_out/r2_52_smoke.log | cut -c1-120
