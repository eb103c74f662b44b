# round 2, GPU call 44: int8 pre-filter for COSINE stores (lower bound from the grid dot product): parity tests, then
# configs[2]'s shape (1 M x 768 cosine) with and without it
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py -x -q -m gpu --timeout 600 -k "prefilter or codec" > gpurun_out/r2_44_pytest.log 2>&1; tail -n 4 gpurun_out/r2_44_pytest.log
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --dims 768 --metric COSINE --nq 10000 --steps 10"
timeout 600 $SB --L 32 64 --tag c3_plain > gpurun_out/r2_44_c3_plain.json 2> gpurun_out/r2_44_c3_plain.err
timeout 600 $SB --L 32 64 --codec --tag c3_codec > gpurun_out/r2_44_c3_codec.json 2> gpurun_out/r2_44_c3_codec.err
grep "'L'" gpurun_out/r2_44_c3_*.err | cut -c1-420
