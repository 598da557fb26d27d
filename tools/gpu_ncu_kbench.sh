#!/bin/bash
# ncu --set full of one kernel inside tools/kbench.py (kernel-only micro-benchmark).
# usage: tools/gpu_ncu_kbench.sh <tag> <kbench-which> <kernel-regex> [env assignments...]
TAG=$1; WHICH=$2; RE=$3; shift 3
mkdir -p gpurun_out
env "$@" timeout 300 python tools/kbench.py $WHICH > gpurun_out/${TAG}_plain.log 2>&1 &&
env "$@" timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RE" -s 4 -c 1 \
  -o /tmp/${TAG}_prof -f python tools/kbench.py $WHICH --reps 3 > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/${TAG}_plain.log
ncu -i /tmp/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ncu -i /tmp/${TAG}_prof.ncu-rep --page source --csv > gpurun_out/${TAG}_source.csv 2>/dev/null
