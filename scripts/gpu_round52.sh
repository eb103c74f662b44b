# round 52: the whole GPU suite on the final code
set -x
timeout 600 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu_r52.log 2>&1; tail -n 4 gpurun_out/pytest_gpu_r52.log
