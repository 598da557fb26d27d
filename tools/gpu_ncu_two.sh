#!/bin/bash
# ncu --set full of the two roofline kernels only (same bench command as gpu_round.sh)
TAG=${1:-x}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
for K in corr_lookup_v2_kernel:2:2 convgru_fused_kernel:2:1; do
  RE=${K%%:*}; R=${K#*:}; CNT=${R%%:*}; SKIP=${R#*:}
  SCFLOW_PROFILE=1 timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:"$RE" -s $SKIP -c $CNT -o /tmp/${TAG}_${RE} -f $CMD > gpurun_out/${TAG}_ncu_${RE}.log 2>&1
  echo "ncu $RE rc=$?"
  ncu -i /tmp/${TAG}_${RE}.ncu-rep --page raw --csv > gpurun_out/${TAG}_${RE}_raw.csv 2>/dev/null
done
