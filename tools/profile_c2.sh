#!/bin/bash
# ncu evidence for the C2 bench (run under gpurun; one GPU).  Outputs under gpurun_out/.
#   1. launch list (gpu__time_duration.sum per launch) of a short bench run
#   2. one --set full capture of the three GEMM kernels (layer FWD, layer BWD, dW) and the streaming kernels
set -u
mkdir -p gpurun_out
TAG=${1:-r01b}
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench_c2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_tc_layer2|k_tc_dw|k_tc_last|k_tc_first' -s 12 -c 8 \
    -o gpurun_out/${TAG}_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_full.log 2>&1
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2> /dev/null
ls -la gpurun_out
