# round 2, GPU call 27: ncu --set full of the final kernels (bench's pre-filtered kernel, int8 route, ternary 768-d) + launch list
set -x
mkdir -p gpurun_out
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 2 --warmup 1"
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
timeout 600 $NCU -k regex:'traverse_kernel<.*\(int\)3, \(int\)0>' -s 6 -c 1 -o gpurun_out/r2_27_prof_bench_pf python bench.py --steps 3 --warmup 3 --L 56 --no-cpu --extras none > gpurun_out/r2_27_ncu_pf.log 2>&1
timeout 600 $NCU -k regex:'traverse_kernel<.*\(int\)1, \(int\)2, \(int\)0>' -s 1 -c 1 -o gpurun_out/r2_27_prof_int8 $SB --dims 128 --route int8 --L 56 > gpurun_out/r2_27_ncu_int8.log 2>&1
timeout 600 $NCU -k regex:'traverse_kernel<.*\(int\)2, \(int\)1, \(int\)0>' -s 1 -c 1 -o gpurun_out/r2_27_prof_tern768 $SB --dims 768 --metric COSINE --route ternary --L 80 > gpurun_out/r2_27_ncu_tern768.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'traverse_kernel|rerank|merge' -c 60 --csv --log-file gpurun_out/r2_27_launches.csv python bench.py --steps 3 --warmup 3 --L 56 --no-cpu --extras none > gpurun_out/r2_27_ncu_launch.log 2>&1
ls -la gpurun_out/r2_27*.ncu-rep
