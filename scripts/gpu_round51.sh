# round 51: filter budget by kernel family (12 / 16 / 24 warps): tests, small batches at L = 128 and L = 56
set -x
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_r51.log 2>&1; tail -n 3 gpurun_out/pytest_gpu_r51.log
timeout 100 python scripts/mode_sweep.py --L 128 --nqs 1 256 1024 1776 2368 > gpurun_out/small_batch_L128_r51.json 2> gpurun_out/small_batch_L128_r51.err
cut -c1-200 gpurun_out/small_batch_L128_r51.err
