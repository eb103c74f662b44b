# round 2, GPU call 3: NaN debug variants, GPU suite, ncu --set full of the code-route / ternary kernels
set -x
mkdir -p gpurun_out
timeout 120 python scripts/debug/special_values.py merge_min=33 > gpurun_out/r2_03_special_insert.log 2>&1; grep -n "row 8" gpurun_out/r2_03_special_insert.log | head -20
timeout 120 python scripts/debug/special_values.py filter_mode=0 > gpurun_out/r2_03_special_f0.log 2>&1; grep -n "row 8" gpurun_out/r2_03_special_f0.log | head -20
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 --deselect tests/test_gpu_round2.py::test_special_values > gpurun_out/r2_03_pytest.log 2>&1; tail -n 15 gpurun_out/r2_03_pytest.log
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 2 --warmup 1"
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
timeout 600 $NCU -k regex:'traverse_kernel<8, 4, 2, 2, 1, 3, 0>' -s 1 -c 1 -o gpurun_out/r2_03_prof_pf $SB --dims 128 --codec --L 56 > gpurun_out/r2_03_ncu_pf.log 2>&1
timeout 600 $NCU -k regex:'traverse_kernel<8, 4, 2, 2, 1, 2, 0>' -s 1 -c 1 -o gpurun_out/r2_03_prof_int8 $SB --dims 128 --route int8 --L 56 > gpurun_out/r2_03_ncu_int8.log 2>&1
timeout 600 $NCU -k regex:'traverse_kernel<.*, 2, 1, 0>' -s 1 -c 1 -o gpurun_out/r2_03_prof_tern768 $SB --dims 768 --metric COSINE --route ternary --L 80 > gpurun_out/r2_03_ncu_tern768.log 2>&1
timeout 600 $NCU -k regex:'rerank_kernel' -s 1 -c 1 -o gpurun_out/r2_03_prof_rerank768 $SB --dims 768 --metric COSINE --route ternary --L 80 > gpurun_out/r2_03_ncu_rerank768.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -n 3 gpurun_out/r2_03_ncu_*.log
