# round 2, GPU call 30: lmd_search_batch_static test alone in a fresh process, then the whole GPU suite
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu --timeout 300 > gpurun_out/r2_30_alone.log 2>&1; tail -n 5 gpurun_out/r2_30_alone.log
timeout 600 python -m pytest "tests/test_zz_no_alloc.py::test_static_scratch_search_equals_search_batch[COSINE-ternary-96]" -q -m gpu --timeout 300 > gpurun_out/r2_30_alone1.log 2>&1; tail -n 3 gpurun_out/r2_30_alone1.log
timeout 2400 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r2_30_pytest.log 2>&1; tail -n 5 gpurun_out/r2_30_pytest.log
