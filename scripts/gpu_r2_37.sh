# round 2, GPU call 37: configs[3] (10 M x 96, 100 k queries, L = 96) with the int8 pre-filter on / off
set -x
mkdir -p gpurun_out
SB="python scripts/shard_bench.py --mode replicated --rows 10000000 --dims 96 --nq 100000 --steps 8 --warmup 2"
timeout 600 $SB --L 96 48 --tag c4_plain > gpurun_out/r2_37_c4_plain.json 2> gpurun_out/r2_37_c4_plain.err
timeout 600 $SB --L 96 48 --codec --tag c4_codec > gpurun_out/r2_37_c4_codec.json 2> gpurun_out/r2_37_c4_codec.err
grep "'L'" gpurun_out/r2_37_c4_*.err | cut -c1-400
