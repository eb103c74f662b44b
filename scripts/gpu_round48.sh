# round 48: filtered search with the 24-warp filter budget up to L = 128; tests
set -x
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_r48.log 2>&1; tail -n 3 gpurun_out/pytest_gpu_r48.log
timeout 200 python scripts/filtered_bench.py > gpurun_out/filtered_r48.json 2> gpurun_out/filtered_r48.err
cat gpurun_out/filtered_r48.err | cut -c1-330
