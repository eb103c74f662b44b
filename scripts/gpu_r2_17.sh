# round 2, GPU call 17: neighbour-row prefetch for admitted candidates (latency), full-size filtered + ternary test, bench --route int8
set -x
mkdir -p gpurun_out
timeout 600 python scripts/latency_probe.py > gpurun_out/r2_17_latency.json 2> gpurun_out/r2_17_latency.err; tail -9 gpurun_out/r2_17_latency.err
timeout 300 python scripts/pipeline_probe.py --nq 1250 --steps 200 --tune nbr_prefetch=0 > gpurun_out/r2_17_pipe_1250_nbp0.json 2> gpurun_out/r2_17_pipe_1250_nbp0.err; tail -n 4 gpurun_out/r2_17_pipe_1250_nbp0.err
timeout 300 python scripts/pipeline_probe.py --nq 1250 --steps 200 --tune nbr_prefetch=1 > gpurun_out/r2_17_pipe_1250_nbp1.json 2> gpurun_out/r2_17_pipe_1250_nbp1.err; tail -n 4 gpurun_out/r2_17_pipe_1250_nbp1.err
timeout 600 python -m pytest tests/test_gpu_fullsize.py -q -m gpu --timeout 600 -k "filtered_ternary" > gpurun_out/r2_17_pytest_fullsize.log 2>&1; tail -n 5 gpurun_out/r2_17_pytest_fullsize.log | cut -c1-300
timeout 600 python bench.py --route int8 --steps 20 --warmup 5 --extras none > gpurun_out/r2_17_bench_int8.json 2> gpurun_out/r2_17_bench_int8.err; tail -n 3 gpurun_out/r2_17_bench_int8.err | cut -c1-300; cut -c1-900 gpurun_out/r2_17_bench_int8.json
