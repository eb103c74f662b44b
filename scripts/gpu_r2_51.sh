# round 2, GPU call 51: the modulo-mapped visited filter under the FILTERED kernels (row f4): 3 072 entries against the default
set -x
mkdir -p gpurun_out
FB="python scripts/filtered_bench.py --fracs 1.0 0.1 --Ls 64 128"
timeout 200 $FB > gpurun_out/r2_51_filt_default.json 2> gpurun_out/r2_51_filt_default.err
timeout 200 $FB --tune filter_mode=3 filter_entries=3072 > gpurun_out/r2_51_filt_m3.json 2> gpurun_out/r2_51_filt_m3.err
grep traverse_ms gpurun_out/r2_51_filt_*.err | cut -c1-330
