set -x
mkdir -p gpurun_out
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout 300 $SB --dims 128 --codec --L 56 --sweep merge_min=2,3,5,8,33 --tag pf_mm > gpurun_out/r2_23_pf_mm.json 2> gpurun_out/r2_23_pf_mm.err
timeout 300 $SB --dims 128 --route int8 --L 56 --sweep merge_min=2,3,5,8,33 --tag int8_mm > gpurun_out/r2_23_int8_mm.json 2> gpurun_out/r2_23_int8_mm.err
grep -o "'merge_min': [0-9]*}, 'ms_per_step': [0-9.]*\|'traverse_ms': [0-9.]*\|'merge_fallbacks_per_q': [0-9.]*" gpurun_out/r2_23_pf_mm.err gpurun_out/r2_23_int8_mm.err | paste - - - 
