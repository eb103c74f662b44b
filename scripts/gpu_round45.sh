# round 45: after the 24-warp filter budget and the 96-d G = 4 rule: tests, headline, auto-mode sweeps
set -x
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_r45.log 2>&1; tail -n 3 gpurun_out/pytest_gpu_r45.log
timeout 300 python bench.py --steps 50 --warmup 5 > gpurun_out/bench_r45.json 2> gpurun_out/bench_r45.err; tail -n 1 gpurun_out/bench_r45.json
timeout 120 python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 128 --nq 10000 --L 56 64 96 128 160 192 256 --steps 10 --tag auto_d128 > gpurun_out/auto_d128_r45.json 2> gpurun_out/auto_d128_r45.err
timeout 120 python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 96 --nq 10000 --L 48 64 96 128 160 256 --steps 10 --tag auto_d96 > gpurun_out/auto_d96_r45.json 2> gpurun_out/auto_d96_r45.err
timeout 200 python scripts/shard_bench.py --mode replicated --rows 10000000 --dims 96 --nq 12500 --L 48 64 96 128 --steps 20 --tag auto_10m_d96 > gpurun_out/auto_10m_d96_r45.json 2> gpurun_out/auto_10m_d96_r45.err
