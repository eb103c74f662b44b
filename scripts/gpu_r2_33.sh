# round 2, GPU call 33: ncu --set full of the pre-filtered kernel with its new default filter (7 CTAs per SM); the same
# filter under the plain exact kernels with two batches in flight
set -x
mkdir -p gpurun_out
for t in "" "--tune filter_mode=3 filter_entries=3504" "--tune filter_mode=3 filter_entries=3000" "--tune filter_mode=2"; do
  n=$(echo "$t" | tr -c 'a-z0-9\n' '_')
  timeout 300 python scripts/pipeline_probe.py --L 56 $t > gpurun_out/r2_33_pipe_plain$n.json 2> gpurun_out/r2_33_pipe_plain$n.err
done
timeout 300 python scripts/pipeline_probe.py --dims 96 --L 64 > gpurun_out/r2_33_pipe_plain96.json 2> gpurun_out/r2_33_pipe_plain96.err
timeout 300 python scripts/pipeline_probe.py --dims 96 --L 64 --tune filter_mode=3 filter_entries=3504 > gpurun_out/r2_33_pipe_plain96_m3.json 2> gpurun_out/r2_33_pipe_plain96_m3.err
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
timeout 600 $NCU -k regex:'traverse_kernel<.*\(int\)3, \(int\)0>' -s 6 -c 1 -o gpurun_out/r2_33_prof_bench_pf python bench.py --steps 3 --warmup 3 --L 56 --no-cpu --extras none > gpurun_out/r2_33_ncu_pf.log 2>&1
ls -la gpurun_out/*.ncu-rep
