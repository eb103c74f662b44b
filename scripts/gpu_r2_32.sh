# round 2, GPU call 32: the modulo-mapped filter as the pre-filtered kernel's default (7 CTAs per SM): suite, then A/B
set -x
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r2_32_pytest.log 2>&1; tail -n 3 gpurun_out/r2_32_pytest.log
for L in 56 100 128; do
  timeout 300 python scripts/pipeline_probe.py --codec --L $L > gpurun_out/r2_32_pipe_auto_L$L.json 2> gpurun_out/r2_32_pipe_auto_L$L.err
  timeout 300 python scripts/pipeline_probe.py --codec --L $L --tune filter_mode=2 > gpurun_out/r2_32_pipe_m2_L$L.json 2> gpurun_out/r2_32_pipe_m2_L$L.err
done
timeout 300 python scripts/pipeline_probe.py --codec --dims 96 --L 64 > gpurun_out/r2_32_pipe_auto_d96.json 2> gpurun_out/r2_32_pipe_auto_d96.err
timeout 300 python scripts/pipeline_probe.py --codec --dims 96 --L 64 --tune filter_mode=2 > gpurun_out/r2_32_pipe_m2_d96.json 2> gpurun_out/r2_32_pipe_m2_d96.err
timeout 900 python bench.py > gpurun_out/r2_32_bench.json 2> gpurun_out/r2_32_bench.err; tail -c 1500 gpurun_out/r2_32_bench.json
