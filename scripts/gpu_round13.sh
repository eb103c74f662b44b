# round 13: shard_bench smoke at 2 GPUs (both modes, small)
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR scripts/shard_bench.py --mode sharded --rows 2000000 --dims 96 --nq 10000 --L 32 64 --steps 5 --tag c5_small > gpurun_out/shard_c5_small_n2.json 2> gpurun_out/shard_c5_small_n2.err
timeout 300 $TR scripts/shard_bench.py --mode replicated --rows 1000000 --dims 96 --nq 10000 --L 32 64 --steps 5 --tag c4_small > gpurun_out/shard_c4_small_n2.json 2> gpurun_out/shard_c4_small_n2.err
tail -n 5 gpurun_out/shard_c5_small_n2.err gpurun_out/shard_c4_small_n2.err
