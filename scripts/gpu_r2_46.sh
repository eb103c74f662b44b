# round 2, GPU call 46: a lone 10k batch: the queries beyond the last full round of resident warps on the small-batch kernels?
set -x
mkdir -p gpurun_out
timeout 600 python scripts/tail_split_probe.py > gpurun_out/r2_46_tail_split.json 2> gpurun_out/r2_46_tail_split.err; grep split gpurun_out/r2_46_tail_split.err
