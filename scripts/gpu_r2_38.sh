# round 2, GPU call 38: bench.py with the int8 code sidecar also on configs[3] (c4) - full default run on one GPU
set -x
mkdir -p gpurun_out
timeout 1200 python bench.py > gpurun_out/r2_38_bench.json 2> gpurun_out/r2_38_bench.err; tail -c 300 gpurun_out/r2_38_bench.json
grep "extra c4" gpurun_out/r2_38_bench.err | cut -c1-1500
