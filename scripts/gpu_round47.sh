# round 47: tests once more on the final dispatch rules; filtered search: visited-filter size
set -x
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_r47.log 2>&1; tail -n 3 gpurun_out/pytest_gpu_r47.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r47.log 2>&1; tail -n 2 gpurun_out/smoke_r47.log
timeout 200 python scripts/filtered_bench.py --fracs 1.0 0.1 --filter-log2 0 11 > gpurun_out/filtered_r47.json 2> gpurun_out/filtered_r47.err
cat gpurun_out/filtered_r47.err
