set -x
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --route ternary --nq 10000 --L 80 --steps 10"
for mm in 2 3 5; do for ts in 32 64; do
timeout 200 $SB --tune merge_min=$mm tie_slots=$ts --tag mm${mm}_ts${ts} 2>&1 | grep "'L'" >> gpurun_out/tune_ternary_r22.log
done; done
