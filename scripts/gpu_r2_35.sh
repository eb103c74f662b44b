# round 2, GPU call 35: suite with the modulo filter rule on the plain 128-d kernels too; CTAs of 1 / 2 / 4 warps with two
# batches in flight (does a faster CTA turnover fill the persistent kernels' tails better?)
set -x
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r2_35_pytest.log 2>&1; tail -n 3 gpurun_out/r2_35_pytest.log
P="python scripts/pipeline_probe.py"
for w in 0 1 2; do
  timeout 300 $P --codec --tune warps_per_cta=$w > gpurun_out/r2_35_pf_w$w.json 2> gpurun_out/r2_35_pf_w$w.err
  timeout 300 $P --tune warps_per_cta=$w > gpurun_out/r2_35_plain_w$w.json 2> gpurun_out/r2_35_plain_w$w.err
  timeout 300 $P --route int8 --tune warps_per_cta=$w > gpurun_out/r2_35_int8_w$w.json 2> gpurun_out/r2_35_int8_w$w.err
  timeout 300 $P --L 64 --tune warps_per_cta=$w > gpurun_out/r2_35_plain64_w$w.json 2> gpurun_out/r2_35_plain64_w$w.err
done
