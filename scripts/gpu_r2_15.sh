# round 2, GPU call 15: single-query latency probe; pre-filtered kernels forced into 64 registers (experimental .so)
set -x
mkdir -p gpurun_out
timeout 600 python scripts/latency_probe.py > gpurun_out/r2_15_latency.json 2> gpurun_out/r2_15_latency.err; cat gpurun_out/r2_15_latency.err | tail -12
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout 300 $SB --dims 128 --codec --L 56 --sweep filter_log2=0,11 --tag pf80 > gpurun_out/r2_15_pf80.json 2> gpurun_out/r2_15_pf80.err
LMD_B200_LIB=$PWD/duckdb-lm-diskann_b200/build_exp/liblmd_b200_pf64.so timeout 300 $SB --dims 128 --codec --L 56 --sweep filter_log2=0,11 --tag pf64 > gpurun_out/r2_15_pf64.json 2> gpurun_out/r2_15_pf64.err
tail -n 2 gpurun_out/r2_15_pf80.err gpurun_out/r2_15_pf64.err | cut -c1-400
timeout 300 python scripts/pipeline_probe.py --codec > gpurun_out/r2_15_pipe_pf80.json 2> gpurun_out/r2_15_pipe_pf80.err; tail -n 4 gpurun_out/r2_15_pipe_pf80.err
LMD_B200_LIB=$PWD/duckdb-lm-diskann_b200/build_exp/liblmd_b200_pf64.so timeout 300 python scripts/pipeline_probe.py --codec > gpurun_out/r2_15_pipe_pf64.json 2> gpurun_out/r2_15_pipe_pf64.err; tail -n 4 gpurun_out/r2_15_pipe_pf64.err
