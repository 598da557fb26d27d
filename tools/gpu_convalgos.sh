#!/bin/bash
# convbench under operand-movement variants + the conv parity tests.  usage: tools/gpu_convalgos.sh <tag> [variants...]
# a variant is a convbench argument string, e.g. "--algo 3" or "--algo 0 --bn 128"
set -u
TAG=$1; shift
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_convtc.py -q -x > gpurun_out/${TAG}_pytest_conv.log 2>&1; echo "pytest conv rc=$?"; tail -3 gpurun_out/${TAG}_pytest_conv.log
I=0
FILES=""
for V in "$@"; do
  I=$((I+1))
  timeout 600 python tools/convbench.py $V > gpurun_out/${TAG}_convbench_$I.txt 2>&1; echo "variant $I [$V] rc=$?"
  FILES="$FILES gpurun_out/${TAG}_convbench_$I.txt"
done
python - $FILES <<'PY'
import sys
cols = [[l for l in open(f)] for f in sys.argv[1:]]
for row in zip(*cols):
    print(row[0][:36] + " | " + " | ".join(r[36:48] for r in row))
PY
