set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -3 > gpurun_out/pytest_gpu_r27.log
timeout 300 python scripts/filtered_bench.py > gpurun_out/filtered_1m_d128_r27.json 2> gpurun_out/filtered_1m_d128_r27.err
timeout 300 python scripts/build_bench.py --n 1000000 --dims 768 --metric COSINE --l-insert 64 --Ls 32 > gpurun_out/build_1m_d768_r27.log 2>&1
