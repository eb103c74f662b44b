set -x
mkdir -p gpurun_out
for extra in "--no-cpu --extras none" "--L 56 --extras none --cpu-seconds 2" "--L 56 --no-cpu"; do timeout 400 python bench.py --steps 5 --warmup 3 $extra 2>&1 >/dev/null | grep "single query"; done
