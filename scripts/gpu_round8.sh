set -x
python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/pytest_gpu_r08.log
timeout 600 python scripts/build_bench.py --n 1000000 --l-insert 64 > gpurun_out/build_1m_L64_r08.log 2>&1
timeout 900 python scripts/build_bench.py --n 1000000 --dims 768 --metric COSINE --l-insert 128 --Ls 32 48 64 96 > gpurun_out/build_1m_d768_r08.log 2>&1
timeout 900 python scripts/build_bench.py --n 10000000 --dims 96 --l-insert 128 --nq 10000 --Ls 32 48 64 96 128 > gpurun_out/build_10m_d96_r08.log 2>&1
