set -x
for G in 3 4 6; do
LMD_B200_LIB=$PWD/duckdb-lm-diskann_b200/liblmd_b200_g$G.so timeout 200 python scripts/mode_sweep.py --dims 768 --metric COSINE --L 48 > gpurun_out/mode_sweep_d768_g$G.json 2> gpurun_out/mode_sweep_d768_g$G.err
done
