set -x
python -X faulthandler -m pytest tests/test_host_adapter.py -m gpu -q -x 2>&1 | tail -60 > gpurun_out/pytest_host_r28.log
