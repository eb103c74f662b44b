set -x
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/pytest_gpu_r04.log; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r04.log
python scripts/tune.py --L 48 56 64 --grid "filter_mode=1;filter_log2=11,12;ctas_per_sm=6,8" > gpurun_out/tune_r04.log 2>&1
python scripts/tune.py --L 96 128 --oracle-sample 0 --grid "filter_mode=1;filter_log2=12,13;ctas_per_sm=0" >> gpurun_out/tune_r04.log 2>&1
python scripts/tune.py --L 192 256 --oracle-sample 0 --grid "filter_mode=1;filter_log2=12,13,14;ctas_per_sm=0" >> gpurun_out/tune_r04.log 2>&1
