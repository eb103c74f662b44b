# round 2, GPU call 14: smoke, N = 1 bench after the small-batch rule, housekeeping ncu captures (prune, filtered, L = 128)
set -x
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2_14_smoke.log 2>&1; tail -n 2 gpurun_out/r2_14_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_14_bench.json 2> gpurun_out/r2_14_bench.err; tail -n 2 gpurun_out/r2_14_bench.err | cut -c1-200; cut -c1-700 gpurun_out/r2_14_bench.json
timeout 300 python bench.py --gpus 1 --single-process --steps 20 --warmup 5 > gpurun_out/r2_14_single_process_n1.json 2> /dev/null; cut -c1-500 gpurun_out/r2_14_single_process_n1.json
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 2 --warmup 1"
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
timeout 600 $NCU -k regex:'prune_kernel' -s 40 -c 1 -o gpurun_out/r2_14_prof_prune $SB --dims 128 --L 56 > gpurun_out/r2_14_ncu_prune.log 2>&1
timeout 600 $NCU -k regex:'traverse_kernel<.*\(int\)4, \(int\)1, \(int\)0, \(int\)0>' -s 1 -c 1 -o gpurun_out/r2_14_prof_L128 $SB --dims 128 --L 128 > gpurun_out/r2_14_ncu_L128.log 2>&1
timeout 600 $NCU -k regex:'traverse_kernel<.*\(int\)0, \(int\)1>' -s 1 -c 1 -o gpurun_out/r2_14_prof_filtered python scripts/filtered_bench.py > gpurun_out/r2_14_ncu_filtered.log 2>&1
ls -la gpurun_out/r2_14*.ncu-rep
