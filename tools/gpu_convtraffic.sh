#!/bin/bash
# DRAM traffic per launch of the top convolution shapes, from `ncu --set full` captures of single launches
# (tools/convbench.py --once) at the bench's own sizes -> gpurun_out/<tag>_conv_traffic.json
# usage: tools/gpu_convtraffic.sh <tag>
set -u
TAG=${1:-x}
mkdir -p gpurun_out
i=0
while IFS='|' read -r LABEL N ONLY; do
  i=$((i+1))
  timeout 300 ncu --set full --clock-control none -k regex:'conv_(patch|tc)_kernel' -c 1 -o /tmp/${TAG}_ct_$i -f \
    python tools/convbench.py --once --n $N --only "$ONLY" > gpurun_out/${TAG}_ct_$i.log 2>&1
  ncu -i /tmp/${TAG}_ct_$i.ncu-rep --page raw --csv > gpurun_out/${TAG}_ct_${i}_raw.csv 2>/dev/null
  echo "$LABEL" > gpurun_out/${TAG}_ct_${i}.label
  echo "ncu $LABEL rc=$?"
done <<'LIST'
3x3/1 64->64 @128x128|128|enc 3x3 64->64 @128 (IN raw+stats)
3x3/1 96->96 @64x64|128|enc 3x3 96->96 @64
3x3/1 128->256 @32x32|64|head 3x3 128->256 @32 + 3x3 ->2
3x3/1 256->192 @32x32|64|mot 3x3 256->192 @32
1x5/1 256->256 @32x32|64|GRU zr epilogue 1x5 256->256 @32
1x5/1 256->128 @32x32|64|GRU gate epilogue 1x5 128+128->128 @32
LIST
python tools/conv_traffic_json.py $TAG
