# round 19: ncu captures - ternary-routing traversal at 768-d (L = 80), exact 768-d kernel with G = 6, headline kernel
set -x
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --nq 10000 --steps 2 --warmup 1"
$SB --route ternary --L 80 > gpurun_out/plain_r19.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'traverse_kernel<.*, 1, 0>' -s 2 -c 1 -o gpurun_out/prof_ternary768_r19 $SB --route ternary --L 80 > gpurun_out/ncu_ternary_r19.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'traverse_kernel<32, 6, 6, 2, 2, 0, 0>' -s 2 -c 1 -o gpurun_out/prof_exact768_r19 $SB --L 48 > gpurun_out/ncu_exact768_r19.log 2>&1
python bench.py --steps 3 --warmup 3 --L 56 --no-cpu > gpurun_out/plain2_r19.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'traverse_kernel|rerank_kernel|merge_topk' -c 40 --csv --log-file gpurun_out/launches_r19.csv python bench.py --steps 3 --warmup 3 --L 56 --no-cpu > gpurun_out/ncu_launch_r19.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:traverse_kernel -s 4 -c 1 -o gpurun_out/prof_traverse_r19 python bench.py --steps 3 --warmup 3 --L 56 --no-cpu > gpurun_out/ncu_full_r19.log 2>&1
ls -la gpurun_out/*.ncu-rep
