set -x
nvidia-smi -L
python -m pytest tests -m gpu -q -x 2>&1 | tail -30 > gpurun_out/pytest_gpu_r02.log; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r02.log
python bench.py --sweep > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err; echo "bench exit $?" >> gpurun_out/bench_r02.err
python bench.py --dims 768 --metric COSINE --sweep --steps 20 --warmup 3 --cpu-seconds 8 > gpurun_out/bench_c3_exact_r02.json 2> gpurun_out/bench_c3_exact_r02.err
python bench.py --dims 768 --metric COSINE --route ternary --sweep --steps 20 --warmup 3 --cpu-seconds 8 > gpurun_out/bench_c3_ternary_r02.json 2> gpurun_out/bench_c3_ternary_r02.err
python bench.py --r 64 --sweep --steps 20 --warmup 3 --no-cpu > gpurun_out/bench_r64_r02.json 2> gpurun_out/bench_r64_r02.err
ls -la gpurun_out
