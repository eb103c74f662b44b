# round 42: 8 x B200, C5 and C4 again with the 96-d kernels that keep 4 rows in flight per team
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 170 $TR scripts/shard_bench.py --mode sharded --rows 100000000 --dims 96 --nq 10000 --L 48 64 96 --steps 20 --tag C5 > gpurun_out/shard_c5_n8_r42.json 2> gpurun_out/shard_c5_n8_r42.err
timeout 150 $TR scripts/shard_bench.py --mode replicated --rows 10000000 --dims 96 --nq 12500 --L 48 64 96 --steps 20 --tag C4 > gpurun_out/shard_c4_n8_r42.json 2> gpurun_out/shard_c4_n8_r42.err
tail -n 3 gpurun_out/shard_c5_n8_r42.json gpurun_out/shard_c4_n8_r42.json
