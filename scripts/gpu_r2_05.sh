# round 2, GPU call 5: occupancy / list-size experiments on the code-route and ternary kernels (ncu r2_04 says issue-bound)
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_round2.py -q -m gpu --timeout 600 -k special > gpurun_out/r2_05_pytest.log 2>&1; tail -n 3 gpurun_out/r2_05_pytest.log
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout 300 $SB --dims 128 --route int8 --L 56 --sweep tie_slots=64,32,8,0 filter_log2=0,11,10 --tag int8_occ > gpurun_out/r2_05_int8_occ.json 2> gpurun_out/r2_05_int8_occ.err
timeout 300 $SB --dims 128 --codec --L 56 --sweep filter_log2=0,11,10 --tag pf_occ > gpurun_out/r2_05_pf_occ.json 2> gpurun_out/r2_05_pf_occ.err
timeout 400 $SB --dims 768 --metric COSINE --route ternary --L 80 --sweep tie_slots=64,16 filter_log2=0,11,10 --tag tern_occ > gpurun_out/r2_05_tern_occ.json 2> gpurun_out/r2_05_tern_occ.err
tail -n 2 gpurun_out/r2_05_*.err
