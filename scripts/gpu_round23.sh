set -x
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --route ternary --nq 10000 --L 80 --steps 10"
for lg in 10 11 12; do
timeout 200 $SB --tune filter_log2=$lg --tag lg$lg 2>&1 | grep "'L'" >> gpurun_out/tune_ternary_r23.log
done
timeout 200 $SB --tune filter_log2=11 prefetch_planes=0 --tag lg11_nopf 2>&1 | grep "'L'" >> gpurun_out/tune_ternary_r23.log
