#!/bin/bash
# ncu evidence for one bench workload (run under gpurun; one GPU).  usage: profile_wl.sh <tag> <workload> <kernel regex> [skip] [count]
set -u
mkdir -p gpurun_out
TAG=$1; WL=$2; RE=$3; SKIP=${4:-0}; CNT=${5:-8}
python bench.py --workload $WL --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench_${WL}.csv \
    python bench.py --workload $WL --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"$RE" -s $SKIP -c $CNT \
    -o gpurun_out/${TAG}_full -f python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_full.log 2>&1
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2> /dev/null
ls -la gpurun_out | grep ${TAG}
