set -x
python -m pytest tests/test_gpu_parity.py -m gpu -q -k "latency" 2>&1 | grep -E "^E|Error|FAILED|passed|failed" | head -30 > gpurun_out/pytest_latency_r15.log
timeout 300 python scripts/latency_bench.py > gpurun_out/latency_1m_d128_r15.json 2> gpurun_out/latency_1m_d128_r15.err
timeout 300 python scripts/latency_bench.py --dims 768 --metric COSINE --L 48 > gpurun_out/latency_1m_d768_r15.json 2> gpurun_out/latency_1m_d768_r15.err
# visited-filter sizing fix on a store above 8.4 M rows
timeout 300 python scripts/shard_bench.py --mode replicated --rows 10000000 --dims 96 --nq 12500 --L 48 64 96 --steps 20 --tag C4_1gpu_after_filter_fix > gpurun_out/shard_c4_n1_r15.json 2> gpurun_out/shard_c4_n1_r15.err
