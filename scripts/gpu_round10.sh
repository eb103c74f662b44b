# round 10: live mirror (row f2) tests
set -x
python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "apply_and_remove" 2>&1 | tail -40 > gpurun_out/pytest_apply_r10.log
python -m pytest tests/test_host_adapter.py -m gpu -q -x 2>&1 | tail -40 > gpurun_out/pytest_host_r10.log
