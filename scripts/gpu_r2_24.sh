# round 2, GPU call 24: list_insert through registers - GPU suite, timings, latency
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 -x > gpurun_out/r2_24_pytest.log 2>&1; tail -n 4 gpurun_out/r2_24_pytest.log | cut -c1-300
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout 300 $SB --dims 128 --route int8 --L 56 --tag int8 > gpurun_out/r2_24_int8.json 2> gpurun_out/r2_24_int8.err
timeout 400 $SB --dims 768 --metric COSINE --route ternary --L 80 --tag tern > gpurun_out/r2_24_tern.json 2> gpurun_out/r2_24_tern.err
timeout 300 $SB --dims 128 --codec --L 56 128 --tag pf > gpurun_out/r2_24_pf.json 2> gpurun_out/r2_24_pf.err
timeout 300 $SB --dims 128 --L 56 --tag plain > gpurun_out/r2_24_plain.json 2> gpurun_out/r2_24_plain.err
grep -o "'L': [0-9]*\|'traverse_ms': [0-9.]*" gpurun_out/r2_24_int8.err gpurun_out/r2_24_tern.err gpurun_out/r2_24_pf.err gpurun_out/r2_24_plain.err | paste - -
timeout 600 python scripts/latency_probe.py > gpurun_out/r2_24_latency.json 2> gpurun_out/r2_24_latency.err; head -2 gpurun_out/r2_24_latency.err
