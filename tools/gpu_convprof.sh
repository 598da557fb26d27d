#!/bin/bash
# convbench timings + one ncu --set full capture per conv shape.  usage: tools/gpu_convprof.sh <tag> [--only SUBSTR]
set -u
TAG=$1; shift
mkdir -p gpurun_out
timeout 600 python tools/convbench.py "$@" > gpurun_out/${TAG}_convbench.txt 2>&1; echo "convbench rc=$?"
cat gpurun_out/${TAG}_convbench.txt
CMD="python tools/convbench.py --once $*"
timeout 300 $CMD > gpurun_out/${TAG}_once.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:conv_tc -o /tmp/${TAG}_conv -f $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
ncu -i /tmp/${TAG}_conv.ncu-rep --page raw --csv > gpurun_out/${TAG}_conv_raw.csv 2>/dev/null
ls -la /tmp/${TAG}_conv.ncu-rep
