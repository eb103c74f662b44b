# round 2, GPU call 34: smaller CTAs (faster turnover between two batches in flight) for the pre-filtered kernel; the
# modulo-mapped filter under the plain exact kernels at several list sizes
set -x
mkdir -p gpurun_out
P="python scripts/pipeline_probe.py"
timeout 300 $P --codec --tune warps_per_cta=2 filter_mode=3 filter_entries=3376 > gpurun_out/r2_34_pf_w2.json 2> gpurun_out/r2_34_pf_w2.err
timeout 300 $P --codec --tune warps_per_cta=1 filter_mode=3 filter_entries=3120 > gpurun_out/r2_34_pf_w1.json 2> gpurun_out/r2_34_pf_w1.err
timeout 300 $P --codec --tune warps_per_cta=4 > gpurun_out/r2_34_pf_w4.json 2> gpurun_out/r2_34_pf_w4.err
for L in 32 64 100; do
  timeout 300 $P --L $L > gpurun_out/r2_34_plain_L${L}_auto.json 2> gpurun_out/r2_34_plain_L${L}_auto.err
  for e in 2560 3000 3568; do
    timeout 300 $P --L $L --tune filter_mode=3 filter_entries=$e > gpurun_out/r2_34_plain_L${L}_$e.json 2> gpurun_out/r2_34_plain_L${L}_$e.err
  done
done
