# round 50: small batches with long lists: does the visited filter want to be larger when few warps are resident?
set -x
timeout 100 python scripts/mode_sweep.py --L 128 --nqs 1 256 1024 1776 --filter-log2 0 12 13 > gpurun_out/small_batch_L128_r50.json 2> gpurun_out/small_batch_L128_r50.err
cut -c1-200 gpurun_out/small_batch_L128_r50.err
