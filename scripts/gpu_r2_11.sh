# round 2, GPU call 11: bench.py with the codec headline + 2 batches in flight (N = 1), ncu of its pre-filtered kernel
set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_11_bench.json 2> gpurun_out/r2_11_bench.err; tail -n 3 gpurun_out/r2_11_bench.err | cut -c1-300; cut -c1-1500 gpurun_out/r2_11_bench.json
timeout 600 python bench.py --steps 20 --warmup 5 --impl reference > gpurun_out/r2_11_bench_ref.json 2> gpurun_out/r2_11_bench_ref.err; cut -c1-600 gpurun_out/r2_11_bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'traverse_kernel|rerank_kernel|merge' -c 60 --csv --log-file gpurun_out/r2_11_launches.csv python bench.py --steps 3 --warmup 3 --L 56 --no-cpu --extras none > gpurun_out/r2_11_ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'traverse_kernel<.*\(int\)3, \(int\)0>' -s 6 -c 1 -o gpurun_out/r2_11_prof_bench_pf python bench.py --steps 3 --warmup 3 --L 56 --no-cpu --extras none > gpurun_out/r2_11_ncu_full.log 2>&1
ls -la gpurun_out/r2_11*.ncu-rep
