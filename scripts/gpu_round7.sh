set -x
python -m pytest tests/test_gpu_build.py -m gpu -q -x 2>&1 | tail -40 > gpurun_out/pytest_build_r07.log
python -m pytest tests -m gpu -q --deselect tests/test_gpu_build.py 2>&1 | tail -5 > gpurun_out/pytest_gpu_r07.log
timeout 600 python scripts/build_bench.py --n 1000000 --l-insert 64 > gpurun_out/build_1m_L64_r07.log 2>&1
timeout 600 python scripts/build_bench.py --n 1000000 --l-insert 128 > gpurun_out/build_1m_L128_r07.log 2>&1
python bench.py --steps 50 --warmup 5 --cpu-seconds 5 > gpurun_out/bench_r07.json 2> gpurun_out/bench_r07.err
