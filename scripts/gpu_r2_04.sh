# round 2, GPU call 4: special values after the NaN-key fix, ncu --set full of the code-route / ternary kernels
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_round2.py tests/test_host_adapter.py -q -m gpu --timeout 600 > gpurun_out/r2_04_pytest.log 2>&1; tail -n 8 gpurun_out/r2_04_pytest.log
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 2 --warmup 1"
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
timeout 600 $NCU -k regex:'traverse_kernel<.*\(int\)3, \(int\)0>' -s 1 -c 1 -o gpurun_out/r2_04_prof_pf $SB --dims 128 --codec --L 56 > gpurun_out/r2_04_ncu_pf.log 2>&1
timeout 600 $NCU -k regex:'traverse_kernel<.*\(int\)1, \(int\)2, \(int\)0>' -s 1 -c 1 -o gpurun_out/r2_04_prof_int8 $SB --dims 128 --route int8 --L 56 > gpurun_out/r2_04_ncu_int8.log 2>&1
timeout 600 $NCU -k regex:'traverse_kernel<.*\(int\)2, \(int\)1, \(int\)0>' -s 1 -c 1 -o gpurun_out/r2_04_prof_tern768 $SB --dims 768 --metric COSINE --route ternary --L 80 > gpurun_out/r2_04_ncu_tern768.log 2>&1
ls -la gpurun_out/*.ncu-rep
