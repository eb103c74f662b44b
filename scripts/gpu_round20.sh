set -x
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --nq 10000 --steps 2 --warmup 1"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'traverse_kernel<.*\(int\)1, \(int\)0>' -s 2 -c 1 -o gpurun_out/prof_ternary768_r19 $SB --route ternary --L 80 > gpurun_out/ncu_ternary_r19.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'traverse_kernel<\(int\)32, \(int\)6, \(int\)6, \(int\)4, \(int\)2, \(int\)0, \(int\)0>' -s 2 -c 1 -o gpurun_out/prof_exact768_r19 $SB --L 100 > gpurun_out/ncu_exact768_r19.log 2>&1
ls -la gpurun_out/*.ncu-rep
