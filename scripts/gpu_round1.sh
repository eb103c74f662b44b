set -x
nvidia-smi -L
python -m pytest tests -m gpu -q -x 2>&1 | tail -40 > gpurun_out/pytest_gpu.log; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
python bench.py --sweep > gpurun_out/bench_r01.json 2> gpurun_out/bench_r01.err; echo "bench exit $?" >> gpurun_out/bench_r01.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_r01.json 2> gpurun_out/bench_ref_r01.err
python bench.py --steps 3 --warmup 3 --L 64 --no-cpu > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'traverse_kernel|rerank_kernel|merge_topk' -c 40 --csv --log-file gpurun_out/launches_r01.csv python bench.py --steps 3 --warmup 3 --L 64 --no-cpu > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:traverse_kernel -s 4 -c 2 -o gpurun_out/prof_traverse_r01 python bench.py --steps 3 --warmup 3 --L 64 --no-cpu > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
