#!/bin/bash
# ncu evidence for the C2 bench in one arithmetic mode (run under gpurun; one GPU).  Outputs under gpurun_out/.
#   usage: tools/profile_c2.sh <tag> <precision>
#   1. the plain bench line of the same command (no ncu)
#   2. launch list (gpu__time_duration.sum per launch) of a short bench run
#   3. one --set full capture of one step's kernels (un-graphed probe: tests/diag/c2_step_probe.py)
set -u
mkdir -p gpurun_out
TAG=${1:-r02}; PREC=${2:-fp16s}
ARGS="--steps 2 --warmup 3 --no-cpu-baseline --extras none --snr-steps 0 --precision $PREC"
python bench.py $ARGS > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench_c2.csv \
    python bench.py $ARGS > gpurun_out/${TAG}_ncu_launch.log 2>&1
# the probe runs 1 untimed + 1 profiled step without a CUDA graph: skip the first step's 20 launches (pack, first, 5 fwd, last, 5 dW, 5 bwd, colsum, adam), capture the second's
ncu --set full --clock-control none --import-source on -k regex:'k_tc_|k_adam' -s 20 -c 20 \
    -o gpurun_out/${TAG}_full -f python tests/diag/c2_step_probe.py $PREC 1 > gpurun_out/${TAG}_ncu_full.log 2>&1
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2> /dev/null
ls -la gpurun_out | grep ${TAG}
