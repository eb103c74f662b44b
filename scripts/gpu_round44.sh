# round 44: long lists (L >= 96): kernel family x visited-filter size, 1 M rows at 128-d and 96-d
set -x
timeout 200 python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 128 --nq 10000 --L 96 128 160 192 256 --sweep latency_mode=0,2 filter_log2=11,12,13 --steps 10 --tag long_lists_d128 > gpurun_out/long_lists_d128_r44.json 2> gpurun_out/long_lists_d128_r44.err
timeout 200 python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 96 --nq 10000 --L 96 128 160 192 256 --sweep latency_mode=0,2 filter_log2=11,12,13 --steps 10 --tag long_lists_d96 > gpurun_out/long_lists_d96_r44.json 2> gpurun_out/long_lists_d96_r44.err
tail -n 2 gpurun_out/long_lists_*_r44.err
