# round 2, GPU call 48: ncu --set full + launch list of the final pre-filtered kernel (3 072-entry modulo filter, 28
# one-warp CTAs per SM) as bench.py launches it
set -x
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
timeout 600 $NCU -k regex:'traverse_kernel<.*\(int\)3, \(int\)0>' -s 6 -c 1 -o gpurun_out/r2_48_prof_bench_pf python bench.py --steps 3 --warmup 3 --L 56 --no-cpu --extras none > gpurun_out/r2_48_ncu_pf.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'traverse_kernel|rerank|merge' -c 60 --csv --log-file gpurun_out/r2_48_launches.csv python bench.py --steps 3 --warmup 3 --L 56 --no-cpu --extras none > gpurun_out/r2_48_ncu_launch.log 2>&1
ls -la gpurun_out/*.ncu-rep
