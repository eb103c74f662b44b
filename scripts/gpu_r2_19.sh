# round 2, GPU call 19: direct-mapped visited filter (mode 2) against the bucket filter on the issue-bound kernels
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu --timeout 300 -k "filter_size" > gpurun_out/r2_19_pytest.log 2>&1; tail -n 3 gpurun_out/r2_19_pytest.log
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout 300 $SB --dims 128 --codec --L 56 --sweep filter_mode=1,2 filter_log2=0,11,12 --tag pf_fm > gpurun_out/r2_19_pf_fm.json 2> gpurun_out/r2_19_pf_fm.err
timeout 300 $SB --dims 128 --route int8 --L 56 --sweep filter_mode=1,2 filter_log2=0,11,12 --tag int8_fm > gpurun_out/r2_19_int8_fm.json 2> gpurun_out/r2_19_int8_fm.err
timeout 400 $SB --dims 768 --metric COSINE --route ternary --L 80 --sweep filter_mode=1,2 filter_log2=0,11,12 --tag tern_fm > gpurun_out/r2_19_tern_fm.json 2> gpurun_out/r2_19_tern_fm.err
timeout 300 $SB --dims 128 --L 56 --sweep filter_mode=1,2 filter_log2=0,12 --tag plain_fm > gpurun_out/r2_19_plain_fm.json 2> gpurun_out/r2_19_plain_fm.err
