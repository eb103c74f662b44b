set -x
python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/pytest_gpu_r05.log
python bench.py --sweep > gpurun_out/bench_r05.json 2> gpurun_out/bench_r05.err
L=$(python -c "import json;print(json.load(open('gpurun_out/bench_r05.json'))['config']['L'])")
python scripts/tune.py --L 56 128 192 --oracle-sample 0 --grid "warps_per_cta=4,2,1" > gpurun_out/tune_r05.log 2>&1
python scripts/tune.py --dims 768 --metric COSINE --route ternary --L 80 --grid "prefetch_planes=0,1;tie_slots=-1,32" >> gpurun_out/tune_r05.log 2>&1
python bench.py --steps 3 --warmup 3 --L $L --no-cpu > gpurun_out/plain_r05.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'traverse_kernel|rerank_kernel|merge_topk' -c 40 --csv --log-file gpurun_out/launches_r05.csv python bench.py --steps 3 --warmup 3 --L $L --no-cpu > gpurun_out/ncu_launch_r05.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:traverse_kernel -s 4 -c 1 -o gpurun_out/prof_traverse_r05 python bench.py --steps 3 --warmup 3 --L $L --no-cpu > gpurun_out/ncu_full_r05.log 2>&1
