# quick GPU check: parity tests + bench with L sweep (no ncu)
python -m pytest tests -m gpu -q -x 2>&1 | tail -30 > gpurun_out/pytest_gpu.log; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python bench.py --sweep --no-cpu > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench exit $?" >> gpurun_out/bench_quick.err
