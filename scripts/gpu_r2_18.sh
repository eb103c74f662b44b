# round 2, GPU call 18: ternary distance with half the popcounts - parity tests and timing
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu --timeout 600 -k "ternary or filtered or sidecar or release" > gpurun_out/r2_18_pytest.log 2>&1; tail -n 5 gpurun_out/r2_18_pytest.log | cut -c1-300
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout 400 $SB --dims 768 --metric COSINE --route ternary --L 80 96 --tag tern768 > gpurun_out/r2_18_tern768.json 2> gpurun_out/r2_18_tern768.err
timeout 400 $SB --dims 128 --metric COSINE --route ternary --L 64 128 --tag tern128 > gpurun_out/r2_18_tern128.json 2> gpurun_out/r2_18_tern128.err
tail -n 2 gpurun_out/r2_18_tern*.err | cut -c1-500
