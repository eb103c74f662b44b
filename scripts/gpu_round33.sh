set -x
python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/pytest_gpu_r33.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r33.log 2>&1
