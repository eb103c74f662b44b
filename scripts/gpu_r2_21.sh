set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/r2_21_pytest.log 2>&1; tail -n 6 gpurun_out/r2_21_pytest.log | cut -c1-300
SB="python scripts/shard_bench.py --mode replicated --rows 1000000 --nq 10000 --steps 10"
timeout 300 $SB --dims 128 --route int8 --L 56 --tag int8 > gpurun_out/r2_21_int8.json 2> gpurun_out/r2_21_int8.err
timeout 400 $SB --dims 768 --metric COSINE --route ternary --L 80 --tag tern > gpurun_out/r2_21_tern.json 2> gpurun_out/r2_21_tern.err
timeout 300 $SB --dims 128 --codec --L 56 --tag pf > gpurun_out/r2_21_pf.json 2> gpurun_out/r2_21_pf.err
tail -n 1 gpurun_out/r2_21_int8.err gpurun_out/r2_21_tern.err gpurun_out/r2_21_pf.err | cut -c1-330
