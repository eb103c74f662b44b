set -x
python -m pytest tests/test_sharded.py -m gpu -q 2>&1 | tail -3 > gpurun_out/pytest_sharded_r26.log
timeout 300 python scripts/filtered_bench.py > gpurun_out/filtered_1m_d128_r26.json 2> gpurun_out/filtered_1m_d128_r26.err
timeout 300 python scripts/build_bench.py --n 1000000 --dims 768 --metric COSINE --l-insert 64 --Ls 32 48 > gpurun_out/build_1m_d768_r26.log 2>&1
timeout 300 python scripts/build_bench.py --n 1000000 --l-insert 64 --Ls 56 > gpurun_out/build_1m_d128_r26.log 2>&1
