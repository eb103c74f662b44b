# round 2, GPU call 9: pre-filtered exact kernels with one row in flight per team (64 registers) x filter size
set -x
mkdir -p gpurun_out
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout 300 $SB --dims 128 --codec --L 56 --sweep pf_g=2,1 filter_log2=0,11 --tag pf_g_d128 > gpurun_out/r2_09_pf_g_d128.json 2> gpurun_out/r2_09_pf_g_d128.err
timeout 300 $SB --dims 128 --codec --L 96 128 --sweep pf_g=2,1 filter_log2=0,11,12 --tag pf_g_d128_long > gpurun_out/r2_09_pf_g_d128_long.json 2> gpurun_out/r2_09_pf_g_d128_long.err
timeout 300 $SB --dims 96 --codec --L 64 --sweep pf_g=2,1 filter_log2=0,11 --tag pf_g_d96 > gpurun_out/r2_09_pf_g_d96.json 2> gpurun_out/r2_09_pf_g_d96.err
tail -n 1 gpurun_out/r2_09_*.err | cut -c1-300
