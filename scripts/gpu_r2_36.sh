# round 2, GPU call 36: final rules (modulo filter: pre-filtered kernel at 28 one-warp CTAs, plain 128-d kernels inside the
# 196 KB carve-out): suite, A/B against the old filters, bench
set -x
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r2_36_pytest.log 2>&1; tail -n 3 gpurun_out/r2_36_pytest.log
P="python scripts/pipeline_probe.py"
timeout 300 $P --codec > gpurun_out/r2_36_pf_auto.json 2> gpurun_out/r2_36_pf_auto.err
timeout 300 $P --codec --dims 96 --L 64 > gpurun_out/r2_36_pf96_auto.json 2> gpurun_out/r2_36_pf96_auto.err
for L in 32 56 64; do
  timeout 300 $P --L $L > gpurun_out/r2_36_plain_L${L}_auto.json 2> gpurun_out/r2_36_plain_L${L}_auto.err
  timeout 300 $P --L $L --tune filter_mode=1 > gpurun_out/r2_36_plain_L${L}_m1.json 2> gpurun_out/r2_36_plain_L${L}_m1.err
done
timeout 900 python bench.py > gpurun_out/r2_36_bench.json 2> gpurun_out/r2_36_bench.err; tail -c 600 gpurun_out/r2_36_bench.json
