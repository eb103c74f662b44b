# round 2, GPU call 41: pre-filtered kernel with 128-entry lists: the 6 KB modulo filter (7 CTAs) against the 4 KB table
set -x
mkdir -p gpurun_out
SB="python scripts/shard_bench.py --mode replicated --rows 10000000 --dims 96 --nq 100000 --steps 8 --warmup 2"
timeout 600 $SB --L 96 --codec --sweep filter_mode=3 filter_entries=2048,3072 --tag c4_m3 > gpurun_out/r2_41_c4_m3.json 2> gpurun_out/r2_41_c4_m3.err
P="python scripts/pipeline_probe.py"
for L in 100 128; do
  timeout 300 $P --codec --L $L --tune filter_mode=3 filter_entries=3072 > gpurun_out/r2_41_pf_L${L}_m3.json 2> gpurun_out/r2_41_pf_L${L}_m3.err
  timeout 300 $P --codec --L $L --tune filter_mode=3 filter_entries=3072 warps_per_cta=1 > gpurun_out/r2_41_pf_L${L}_m3w1.json 2> gpurun_out/r2_41_pf_L${L}_m3w1.err
  timeout 300 $P --codec --L $L > gpurun_out/r2_41_pf_L${L}_auto.json 2> gpurun_out/r2_41_pf_L${L}_auto.err
done
grep "'L'" gpurun_out/r2_41_c4_*.err | cut -c1-330
