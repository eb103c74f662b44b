set -x
python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "ternary or filter" 2>&1 | tail -5 > gpurun_out/pytest_ternary_r24.log
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --route ternary --nq 10000 --L 80 --steps 10"
for grp in 0 1; do for lg in 10 11 12; do
timeout 200 $SB --tune filter_log2=$lg ternary_group=$grp --tag grp${grp}_lg$lg 2>&1 | grep "'L'" | sed "s/^/grp$grp lg$lg /" >> gpurun_out/tune_ternary_r24.log
done; done
