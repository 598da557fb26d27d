#!/bin/bash
# ncu --set full of selected hand-written kernels inside bench.py's timed region.
# usage: tools/gpu_ncu.sh <tag> <kernel-regex> [count] [skip]
set -u
TAG=$1; RE=$2; CNT=${3:-2}; SKIP=${4:-0}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
SCFLOW_PROFILE=1 timeout 600 $CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
SCFLOW_PROFILE=1 timeout 1200 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:"$RE" -s $SKIP -c $CNT -o /tmp/${TAG}_prof -f $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
ncu -i /tmp/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ncu -i /tmp/${TAG}_prof.ncu-rep --page details --csv > gpurun_out/${TAG}_details.csv 2>/dev/null
ncu -i /tmp/${TAG}_prof.ncu-rep --page source --csv --kernel-name regex:"$RE" > gpurun_out/${TAG}_source.csv 2>/dev/null
SZ=$(stat -c %s /tmp/${TAG}_prof.ncu-rep 2>/dev/null || echo 0)
echo "rep size $SZ"
if [ "$SZ" -lt 20000000 ] && [ "$SZ" -gt 0 ]; then cp /tmp/${TAG}_prof.ncu-rep gpurun_out/; fi
ls -la gpurun_out | tail -8
