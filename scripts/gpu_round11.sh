set -x
python -m pytest tests/test_gpu_parity.py -m gpu -q -k "filter" 2>&1 | grep -E "^E|Error|FAILED|passed|failed" | head -40 > gpurun_out/pytest_filtered_r11.log
