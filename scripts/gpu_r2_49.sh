# round 2, GPU call 49: int8 pre-filter for IP stores: its tests first, then the whole GPU suite and smoke
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_round2.py tests/test_host_adapter.py -q -m gpu --timeout 300 -k "prefilter or codec" > gpurun_out/r2_49_pf.log 2>&1; tail -n 6 gpurun_out/r2_49_pf.log
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r2_49_pytest.log 2>&1; tail -n 3 gpurun_out/r2_49_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_49_smoke.log 2>&1; tail -n 1 gpurun_out/r2_49_smoke.log | cut -c1-200
