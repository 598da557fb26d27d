#!/bin/bash
# ncu --set full of selected conv shapes under each operand-movement variant.  usage: tools/gpu_convncu.sh <tag> "<only>" [algos]
set -u
TAG=$1; ONLY=$2; ALGOS=${3:-"1 2 3"}
mkdir -p gpurun_out
for A in $ALGOS; do
  CMD="python tools/convbench.py --once --algo $A --only"
  timeout 300 $CMD "$ONLY" > gpurun_out/${TAG}_once$A.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_ -o /tmp/${TAG}_a$A -f $CMD "$ONLY" > gpurun_out/${TAG}_ncu$A.log 2>&1
  echo "ncu algo $A rc=$?"
  ncu -i /tmp/${TAG}_a$A.ncu-rep --page raw --csv > gpurun_out/${TAG}_a${A}_raw.csv 2>/dev/null
  ncu -i /tmp/${TAG}_a$A.ncu-rep --page details --csv > gpurun_out/${TAG}_a${A}_details.csv 2>/dev/null
  ncu -i /tmp/${TAG}_a$A.ncu-rep --page source --csv > gpurun_out/${TAG}_a${A}_source.csv 2>/dev/null
done
ls -la gpurun_out | grep ${TAG}
