set -x
# C1 (BASELINE.json configs[0]): 10k x 128 L2, R=32, L=64, 1k queries, k=10 - our arm and the reference's CPU arm
python bench.py --n 10000 --nq 1000 --L 64 --steps 50 --warmup 5 --cpu-seconds 5 > gpurun_out/bench_c1_r31.json 2> gpurun_out/bench_c1_r31.err
python bench.py --impl reference --n 10000 --nq 1000 --L 64 --steps 5 --warmup 1 > gpurun_out/bench_c1_ref_r31.json 2> gpurun_out/bench_c1_ref_r31.err
