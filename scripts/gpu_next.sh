# Pending measurements, ready for the next GPU call - about 4 minutes of box time on one B200.
# 1. the whole GPU suite; 2. 10 M x 128 (stores above 8.4 M rows keep the 8 KB filter: which family at L >= 128?);
# 3. 96-d filtered search: G = 2 against the filter sizes; 4. launch list + one full capture of the long-list kernel.
set -x
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu_next.log 2>&1; tail -n 3 gpurun_out/pytest_gpu_next.log
timeout 300 python scripts/shard_bench.py --mode replicated --rows 10000000 --dims 128 --nq 10000 --L 56 96 128 192 \
  --sweep latency_mode=0,2 --steps 10 --tag long_lists_10m_d128 > gpurun_out/long_lists_10m_d128.json 2> gpurun_out/long_lists_10m_d128.err
timeout 200 python scripts/filtered_bench.py --dims 96 --fracs 1.0 0.1 --filter-log2 0 11 12 > gpurun_out/filtered_d96.json 2> gpurun_out/filtered_d96.err
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 128 --nq 10000 --L 128 --steps 3 --warmup 3"
$SB > gpurun_out/plain_L128.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'traverse_kernel' -c 20 --csv \
  --log-file gpurun_out/launches_L128.csv $SB > gpurun_out/ncu_launch_L128.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:traverse_kernel -s 4 -c 1 -o gpurun_out/prof_traverse_L128 \
  $SB > gpurun_out/ncu_full_L128.log 2>&1
ls -la gpurun_out/*.ncu-rep
