set -x
timeout 200 python scripts/mode_sweep.py > gpurun_out/mode_sweep_d128.json 2> gpurun_out/mode_sweep_d128.err
timeout 200 python scripts/mode_sweep.py --dims 96 --L 64 > gpurun_out/mode_sweep_d96.json 2> gpurun_out/mode_sweep_d96.err
timeout 200 python scripts/mode_sweep.py --dims 768 --metric COSINE --L 48 > gpurun_out/mode_sweep_d768.json 2> gpurun_out/mode_sweep_d768.err
