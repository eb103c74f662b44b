# round 2, GPU call 31: modulo-mapped visited filter of any size (mode 3): a 7th CTA per SM for the pre-filtered kernel?
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py -q -m gpu --timeout 300 -k "filter_size or duplicate" > gpurun_out/r2_31_pytest.log 2>&1; tail -n 3 gpurun_out/r2_31_pytest.log
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 20"
timeout 400 $SB --dims 128 --codec --L 56 --sweep filter_mode=3 filter_entries=4096,3504,3072,2560 --tag pf_m3 > gpurun_out/r2_31_pf_m3.json 2> gpurun_out/r2_31_pf_m3.err
timeout 400 $SB --dims 128 --route int8 --L 56 --sweep filter_mode=3 filter_entries=2048,2560,2880,3400,4096 --tag int8_m3 > gpurun_out/r2_31_int8_m3.json 2> gpurun_out/r2_31_int8_m3.err
timeout 400 $SB --dims 768 --metric COSINE --route ternary --L 80 --sweep filter_mode=3 filter_entries=2048,2560,3000,3500,4096 --tag tern_m3 > gpurun_out/r2_31_tern_m3.json 2> gpurun_out/r2_31_tern_m3.err
timeout 400 $SB --dims 128 --L 56 --sweep filter_mode=3 filter_entries=2048,2560,3000,3504,4096 --tag plain_m3 > gpurun_out/r2_31_plain_m3.json 2> gpurun_out/r2_31_plain_m3.err
for e in 3504 3072 2560; do timeout 300 python scripts/pipeline_probe.py --codec --tune filter_mode=3 filter_entries=$e > gpurun_out/r2_31_pipe_m3_$e.json 2> gpurun_out/r2_31_pipe_m3_$e.err; done
timeout 300 python scripts/pipeline_probe.py --codec > gpurun_out/r2_31_pipe_m2.json 2> gpurun_out/r2_31_pipe_m2.err
for e in 2560 3400; do timeout 300 python scripts/pipeline_probe.py --route int8 --tune filter_mode=3 filter_entries=$e > gpurun_out/r2_31_pipe_int8_m3_$e.json 2> gpurun_out/r2_31_pipe_int8_m3_$e.err; done
timeout 300 python scripts/pipeline_probe.py --route int8 > gpurun_out/r2_31_pipe_int8_m2.json 2> gpurun_out/r2_31_pipe_int8_m2.err
