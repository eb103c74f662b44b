set -x
python -m pytest tests/test_gpu_parity.py -m gpu -q -k "latency or exact_routing" 2>&1 | tail -3 > gpurun_out/pytest_latency_r30.log
timeout 300 python scripts/latency_bench.py > gpurun_out/latency_1m_d128_r30.json 2> gpurun_out/latency_1m_d128_r30.err
