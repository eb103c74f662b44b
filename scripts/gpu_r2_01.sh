# round 2, GPU call 1: full GPU suite + first measurements of the routing sidecars (int8 pre-filter / int8 route / ternary planes)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/r2_01_pytest.log 2>&1; tail -n 15 gpurun_out/r2_01_pytest.log
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout 300 $SB --dims 128 --codec --L 56 128 --sweep prefilter=0,1 prefetch_planes=0,1 --tag pf_d128 > gpurun_out/r2_01_pf_d128.json 2> gpurun_out/r2_01_pf_d128.err
timeout 300 $SB --dims 128 --codec --L 56 --sweep code_sigmas=30,50,60 --tag pf_sigmas > gpurun_out/r2_01_pf_sigmas.json 2> gpurun_out/r2_01_pf_sigmas.err
timeout 300 $SB --dims 128 --route int8 --L 48 56 64 80 96 128 --sweep prefetch_planes=0,1 --tag int8_d128 > gpurun_out/r2_01_int8_d128.json 2> gpurun_out/r2_01_int8_d128.err
timeout 400 $SB --dims 768 --metric COSINE --route ternary --L 80 96 --sweep filter_log2=0,10,11,12 prefetch_planes=0,1 --tag tern768 > gpurun_out/r2_01_tern768.json 2> gpurun_out/r2_01_tern768.err
timeout 300 $SB --dims 96 --codec --L 64 --sweep prefilter=0,1 --tag pf_d96 > gpurun_out/r2_01_pf_d96.json 2> gpurun_out/r2_01_pf_d96.err
tail -n 3 gpurun_out/r2_01_*.err
