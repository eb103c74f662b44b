set -x
python -m pytest tests/test_host_adapter.py -m gpu -q 2>&1 | grep -E "^E|Error|FAILED|passed|failed" | head -40 > gpurun_out/pytest_host_r12.log
