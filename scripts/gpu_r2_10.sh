# round 2, GPU call 10: do consecutive batches overlap on different context streams?
set -x
mkdir -p gpurun_out
timeout 300 python scripts/pipeline_probe.py > gpurun_out/r2_10_pipe_plain.json 2> gpurun_out/r2_10_pipe_plain.err; tail -n 4 gpurun_out/r2_10_pipe_plain.err
timeout 300 python scripts/pipeline_probe.py --codec > gpurun_out/r2_10_pipe_pf.json 2> gpurun_out/r2_10_pipe_pf.err; tail -n 4 gpurun_out/r2_10_pipe_pf.err
timeout 300 python scripts/pipeline_probe.py --route int8 > gpurun_out/r2_10_pipe_int8.json 2> gpurun_out/r2_10_pipe_int8.err; tail -n 4 gpurun_out/r2_10_pipe_int8.err
timeout 300 python scripts/pipeline_probe.py --nq 1250 --steps 200 > gpurun_out/r2_10_pipe_1250.json 2> gpurun_out/r2_10_pipe_1250.err; tail -n 4 gpurun_out/r2_10_pipe_1250.err
