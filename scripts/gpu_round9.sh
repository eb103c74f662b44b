# round 9 (re-entry check): everything the driver runs at round end, on the rebuilt library
set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/pytest_gpu_r09.log; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r09.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r09.log 2>&1
python bench.py > gpurun_out/bench_r09.json 2> gpurun_out/bench_r09.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_r09.json 2> gpurun_out/bench_ref_r09.err
