set -x
python -m pytest tests/test_gpu_parity.py -m gpu -q -k "latency" 2>&1 | tail -2 > gpurun_out/pytest_latency_r40.log
timeout 200 python scripts/mode_sweep.py > gpurun_out/mode_sweep_d128_r40.json 2> gpurun_out/mode_sweep_d128_r40.err
timeout 200 python scripts/mode_sweep.py --dims 96 --L 64 > gpurun_out/mode_sweep_d96_r40.json 2> gpurun_out/mode_sweep_d96_r40.err
