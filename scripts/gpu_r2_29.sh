# round 2, GPU call 29: lmd_search_batch_static (kernel scratch in a static device array): its test alone in a fresh
# process, then the whole GPU suite (the new file is collected last), then smoke
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu --timeout 300 > gpurun_out/r2_29_alone.log 2>&1; tail -n 5 gpurun_out/r2_29_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r2_29_pytest.log 2>&1; tail -n 5 gpurun_out/r2_29_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_29_smoke.log 2>&1; tail -n 2 gpurun_out/r2_29_smoke.log
