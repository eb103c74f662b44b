set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -3 > gpurun_out/pytest_gpu_r41.log
timeout 300 python scripts/shard_bench.py --mode replicated --rows 10000000 --dims 96 --nq 12500 --L 48 64 96 --steps 20 --tag C4_1gpu_G4 > gpurun_out/shard_c4_n1_r41.json 2> gpurun_out/shard_c4_n1_r41.err
timeout 300 python scripts/latency_bench.py > gpurun_out/latency_1m_d128_r41.json 2> gpurun_out/latency_1m_d128_r41.err
