# final check of the tree as committed: what the driver runs at round end
set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > gpurun_out/pytest_gpu_r36.log; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r36.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r36.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_r36.json 2> gpurun_out/bench_ref_r36.err
python bench.py > gpurun_out/bench_r36.json 2> gpurun_out/bench_r36.err
