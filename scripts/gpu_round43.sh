# round 43: 96-d rows, G = 2 (latency_mode 0) against G = 4 (latency_mode 2) per list size, 10 M and 1 M rows
set -x
timeout 200 python scripts/shard_bench.py --mode replicated --rows 10000000 --dims 96 --nq 12500 --L 48 64 80 96 128 160 --latency-modes 0 2 --steps 20 --tag g_sweep_10m > gpurun_out/g_sweep_10m_r43.json 2> gpurun_out/g_sweep_10m_r43.err
timeout 120 python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 96 --nq 10000 --L 48 64 80 96 128 160 --latency-modes 0 2 --steps 20 --tag g_sweep_1m > gpurun_out/g_sweep_1m_r43.json 2> gpurun_out/g_sweep_1m_r43.err
timeout 120 python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 128 --nq 10000 --L 56 64 96 128 160 --latency-modes 0 2 --steps 20 --tag g_sweep_1m_d128 > gpurun_out/g_sweep_1m_d128_r43.json 2> gpurun_out/g_sweep_1m_d128_r43.err
tail -n 2 gpurun_out/g_sweep_*_r43.err
