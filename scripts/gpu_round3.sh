set -x
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/pytest_gpu_r03.log; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r03.log
python scripts/tune.py --L 64 128 256 --grid "filter_mode=0,1;filter_log2=0,11,12,13,14" > gpurun_out/tune_filter_r03.log 2>&1
python scripts/tune.py --L 64 --oracle-sample 0 --grid "filter_mode=1;filter_log2=12;ctas_per_sm=4,5,6,7,8,9,10" >> gpurun_out/tune_filter_r03.log 2>&1
