# round 2, GPU call 50: bench.py (no flags) on the final code
set -x
mkdir -p gpurun_out
timeout 1200 python bench.py > gpurun_out/r2_50_bench.json 2> gpurun_out/r2_50_bench.err; tail -c 300 gpurun_out/r2_50_bench.json
