#!/bin/bash
# convbench under each operand-movement variant + the conv parity tests.  usage: tools/gpu_convalgos.sh <tag>
set -u
TAG=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_convtc.py -q -x > gpurun_out/${TAG}_pytest_conv.log 2>&1; echo "pytest conv rc=$?"; tail -3 gpurun_out/${TAG}_pytest_conv.log
for A in 1 2 3; do
  timeout 600 python tools/convbench.py --algo $A > gpurun_out/${TAG}_convbench_algo$A.txt 2>&1; echo "algo $A rc=$?"
done
paste -d'|' <(cut -c1-48 gpurun_out/${TAG}_convbench_algo1.txt) <(cut -c37-48 gpurun_out/${TAG}_convbench_algo2.txt) <(cut -c37-48 gpurun_out/${TAG}_convbench_algo3.txt)
